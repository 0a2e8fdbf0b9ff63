"""TEST INFRASTRUCTURE: builds tests/_cpp/b200_cs_parity (the header-class parity driver) against the reference
headers where they lie under /root/reference.  The binary is git-ignored but travels to the GPU box."""
import os
import subprocess

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
OUT = os.path.join(ROOT, "tests", "_cpp")
BIN = os.path.join(OUT, "b200_cs_parity")
REF = "/root/reference"
GMP = "/lib/x86_64-linux-gnu/libgmp.so.10"


def _flags():
    return ["g++", "-std=c++20", "-O2", "-march=x86-64-v3", "-fopenmp", "-DOPENMP", "-w", "-I" + os.path.join(ROOT, "oracle", "shim")]


def _tail():
    return ["-L" + os.path.join(ROOT, "cofhe_b200"), "-lcofhe_b200", GMP, "-lcrypto", "-Wl,-rpath,$ORIGIN/../../cofhe_b200"]


LOCAL_B200 = os.path.join(OUT, "local_b200")
LOCAL_CPU = os.path.join(OUT, "local_cpu")


def build_local(force=False):
    """The reference's benchmarks/local.cpp, UNMODIFIED (compiled where it lies), twice: on the device backend (this
    repository's include/ first on the path and -DCOFHE_B200: include/cofhe.hpp) and on the reference class."""
    if not os.path.isdir(REF):
        return None
    src = REF + "/benchmarks/local.cpp"
    deps = [src, os.path.join(ROOT, "include", "cofhe.hpp"), os.path.join(ROOT, "include", "cuda", "cuda_cryptosystem.hpp"),
            os.path.join(ROOT, "include", "cofhe_b200.h")]
    os.makedirs(OUT, exist_ok=True)
    if force or not os.path.exists(LOCAL_B200) or any(os.path.getmtime(d) > os.path.getmtime(LOCAL_B200) for d in deps):
        subprocess.check_call(_flags() + ["-DCOFHE_B200", "-I" + os.path.join(ROOT, "include"), "-I" + REF + "/include",
                                          "-I" + REF + "/thirdparty/bicycl/src", src, "-o", LOCAL_B200] + _tail())
    if force or not os.path.exists(LOCAL_CPU):
        subprocess.check_call(_flags() + ["-I" + REF + "/include", "-I" + REF + "/thirdparty/bicycl/src", src, "-o", LOCAL_CPU,
                                          GMP, "-lcrypto"])
    return LOCAL_B200


def build(force=False):
    if not os.path.isdir(REF):
        return None
    build_local(force)
    src = os.path.join(HERE, "b200_cs_parity.cpp")
    deps = [src, os.path.join(ROOT, "include", "cuda", "cuda_cryptosystem.hpp"), os.path.join(ROOT, "include", "cofhe_b200.h")]
    if not force and os.path.exists(BIN) and all(os.path.getmtime(d) < os.path.getmtime(BIN) for d in deps):
        return BIN
    os.makedirs(OUT, exist_ok=True)
    cmd = ["g++", "-std=c++20", "-O2", "-march=x86-64-v3", "-fopenmp", "-DOPENMP", "-w",
           "-I" + os.path.join(ROOT, "oracle", "shim"), "-I" + os.path.join(ROOT, "include"), "-I" + REF + "/include",
           "-I" + REF + "/thirdparty/bicycl/src", src, "-o", BIN, "-L" + os.path.join(ROOT, "cofhe_b200"), "-lcofhe_b200",
           GMP, "-lcrypto", "-Wl,-rpath,$ORIGIN/../../cofhe_b200"]
    subprocess.check_call(cmd)
    return BIN


if __name__ == "__main__":
    print(build(force=True), LOCAL_B200, LOCAL_CPU)
