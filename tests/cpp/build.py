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


def build(force=False):
    if not os.path.isdir(REF):
        return None
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
    print(build(force=True))
