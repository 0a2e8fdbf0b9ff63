"""The drop-in C++ class (include/cuda/cuda_cryptosystem.hpp): compiles against the reference's headers with
CryptoSystem = B200CryptoSystem (CPU, where /root/reference exists), and on the GPU gives the reference's results."""
import os
import subprocess
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "tests", "cpp"))
BIN = os.path.join(ROOT, "tests", "_cpp", "b200_cs_parity")


def test_header_class_compiles():
    if not os.path.isdir("/root/reference"):
        pytest.skip("reference headers not present on this machine")
    from cofhe_b200 import build as B
    B.build()
    import build as cppbuild
    assert cppbuild.build() == BIN and os.path.exists(BIN)


@pytest.mark.gpu
@pytest.mark.parametrize("nbits,k", [(128, 128), (64, 16), (128, 256)])
def test_header_class_matches_reference(nbits, k):
    if not os.path.exists(BIN):
        pytest.skip("tests/_cpp/b200_cs_parity was not built (needs /root/reference at build time)")
    r = subprocess.run([BIN, str(nbits), str(k)], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "mismatches 0" in r.stdout, r.stdout + r.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("nbits,k", [(128, 128), (64, 16), (256, 64)])
def test_header_class_compact_variant_matches_reference(nbits, k):
    """CPUCryptoSystem(level, k, compact=true) (CL_HSM2k.inl:31-39, 76-80, 458-474): every call of the parity driver --
    encrypt, add, scal / matmul, negate, decrypt, threshold decryption, the single-ciphertext calls, the wire formats and
    the device-resident chain -- with c1, h and pk in Cl(DeltaK), against the reference class draw for draw"""
    if not os.path.exists(BIN):
        pytest.skip("tests/_cpp/b200_cs_parity was not built (needs /root/reference at build time)")
    r = subprocess.run([BIN, str(nbits), str(k), "1"], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0 and "mismatches 0" in r.stdout, r.stdout + r.stderr


LOCAL_B200 = os.path.join(ROOT, "tests", "_cpp", "local_b200")
LOCAL_CPU = os.path.join(ROOT, "tests", "_cpp", "local_cpu")


def test_reference_benchmark_builds_unchanged_on_backend():
    """benchmarks/local.cpp, compiled where it lies under /root/reference with only flags added (-DCOFHE_B200, this
    repository's include/ first), links against the C ABI: make_cryptosystem and the name CPUCryptoSystem resolve to
    the device backend (include/cofhe.hpp, include/cuda/cuda_cryptosystem.hpp)."""
    if not os.path.isdir("/root/reference"):
        pytest.skip("reference sources not present on this machine")
    import build as cppbuild
    assert cppbuild.build_local() == LOCAL_B200
    syms = subprocess.run(["nm", "-D", LOCAL_B200], capture_output=True, text=True).stdout
    for s in ("cofhe_b200_create", "cofhe_b200_add_ciphertexts", "cofhe_b200_scal_matmul", "cofhe_b200_encrypt"):
        assert s in syms
    assert "cofhe_b200_" not in subprocess.run(["nm", "-D", LOCAL_CPU], capture_output=True, text=True).stdout


@pytest.mark.gpu
@pytest.mark.parametrize("mode", ["ciphertext_matadd", "scal_matmul"])
def test_reference_benchmark_runs_unchanged_on_backend(mode, tmp_path):
    """the unmodified reference driver runs both benchmark modes to completion on the device backend (it prints timings only;
    the bit-for-bit comparison of the same chained calls against the reference class is test_class_bench_parity)"""
    if not os.path.exists(LOCAL_B200):
        pytest.skip("tests/_cpp/local_b200 was not built (needs /root/reference at build time)")
    r = subprocess.run([LOCAL_B200, mode], capture_output=True, text=True, timeout=900, cwd=tmp_path)
    assert r.returncode == 0, r.stdout + r.stderr
    assert "n: " in r.stdout


@pytest.mark.gpu
@pytest.mark.parametrize("args", [["matadd", "64", "64", "50", "50"], ["matmul", "8", "64", "64", "50", "2"]])
def test_class_bench_parity(args):
    """the chained calls of benchmarks/local.cpp (ciphertext_matadd 64x64 x 50; scal_matmul 8x64 . 64x64 x 50) through
    B200CryptoSystem on Tensor<CipherText*>: tensors equal the reference class's bit for bit"""
    if not os.path.exists(BIN):
        pytest.skip("tests/_cpp/b200_cs_parity was not built")
    import json
    r = subprocess.run([BIN, "--bench"] + args, capture_output=True, text=True, timeout=1200)
    assert r.returncode == 0, r.stdout + r.stderr
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["mismatches"] == 0 and line["launches"] > 0


@pytest.mark.gpu
def test_class_shared_by_threads():
    """ONE B200CryptoSystem shared by 8 host threads, as the reference's server shares its cryptosystem between its worker
    threads (include/node/server.hpp:16,187-196): concurrent chained add_ciphertext_tensors + part_decrypt_tensor calls give
    the reference's results, and run on the object's concurrency lanes instead of queueing on one stream"""
    if not os.path.exists(BIN):
        pytest.skip("tests/_cpp/b200_cs_parity was not built")
    import json
    r = subprocess.run([BIN, "--threads", "8", "32", "32", "6"], capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout + r.stderr
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["mismatches"] == 0 and line["lanes"] >= 2
    env = dict(os.environ, COFHE_B200_LANES="1")
    r = subprocess.run([BIN, "--threads", "4", "16", "16", "3"], capture_output=True, text=True, timeout=900, env=env)
    assert r.returncode == 0, r.stdout + r.stderr
    line = json.loads(r.stdout.strip().splitlines()[-1])
    assert line["mismatches"] == 0 and line["lanes"] == 1
