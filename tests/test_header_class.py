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
