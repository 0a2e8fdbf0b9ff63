"""Builds libcofhe_b200.so (CUDA, sm_100a) in-tree with nvcc.  Used by __graft_entry__.build() and by hand:
    python -m cofhe_b200.build [--force] [-v]
Every .cu under csrc/ is one translation unit (kernels.cu: C ABI + generic path; fast_n<N>.cu: the register-resident
kernels of one limb count); stale units are compiled in parallel into csrc/_obj/ and linked.  The library and the
objects are git-ignored; the library travels to the GPU box with the working tree.
COFHE_B200_NVCC_EXTRA="-DCOFHE_B200_DEV_N7" builds only the T0 (7-limb) fast kernels while iterating."""
import os
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(HERE)
CSRC = os.path.join(HERE, "csrc")
OBJ = os.path.join(CSRC, "_obj")
LIB = os.path.join(HERE, "libcofhe_b200.so")
NVCC = os.environ.get("NVCC", "/usr/local/cuda/bin/nvcc")
FLAGS = ["-gencode", "arch=compute_100a,code=sm_100a", "-lineinfo", "-O3", "-std=c++17", "-Xcompiler", "-fPIC",
         "-I" + os.path.join(ROOT, "include"), "-I" + CSRC]


def extra_flags():
    return os.environ.get("COFHE_B200_NVCC_EXTRA", "").split()


def sources():
    names = sorted(f for f in os.listdir(CSRC) if f.endswith(".cu"))
    if "-DCOFHE_B200_DEV_N7" in extra_flags():
        names = [f for f in names if not f.startswith(("fast_n", "coop_n")) or f in ("fast_n7.cu", "coop_n7.cu")]
    return [os.path.join(CSRC, f) for f in names]


def dep_files():
    return sorted([os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".cu"))] + [os.path.join(ROOT, "include", "cofhe_b200.h")])


def deps_mtime():
    return max(os.path.getmtime(d) for d in dep_files())


def source_hash():
    """Content hash of everything the library is built from (sources, headers, flags).  Staleness is decided by content,
    not by mtime: the working tree is copied to the GPU box, where the prebuilt library must be taken as it is."""
    import hashlib
    h = hashlib.sha256(" ".join(FLAGS[:6] + extra_flags()).encode())
    for d in dep_files():
        h.update(os.path.basename(d).encode())
        with open(d, "rb") as fh:
            h.update(fh.read())
    return h.hexdigest()


STAMP = LIB + ".srchash"


def obj_of(src):
    return os.path.join(OBJ, os.path.basename(src)[:-3] + ".o")


def stale():
    if not os.path.exists(LIB) or not os.path.exists(STAMP):
        return True
    with open(STAMP) as fh:
        return fh.read().strip() != source_hash()


def build(force=False, verbose=False):
    if not force and not stale():
        return LIB
    os.makedirs(OBJ, exist_ok=True)
    tag = os.path.join(OBJ, "flags.txt")
    flags = " ".join(FLAGS + extra_flags())
    if not os.path.exists(tag) or open(tag).read() != flags:
        force = True
    newest = deps_mtime()
    procs = []
    for src in sources():
        obj = obj_of(src)
        # headers are shared by every unit: a unit is rebuilt when anything under csrc/ is newer than its object
        if force or not os.path.exists(obj) or os.path.getmtime(obj) < newest:
            cmd = [NVCC] + FLAGS + extra_flags() + (["-Xptxas", "-v"] if verbose else []) + ["-c", "-o", obj, src]
            procs.append((src, subprocess.Popen(cmd)))
    failed = [src for src, p in procs if p.wait() != 0]
    if failed:
        raise RuntimeError("nvcc failed for " + ", ".join(os.path.basename(f) for f in failed))
    with open(tag, "w") as fh:
        fh.write(flags)
    subprocess.check_call([NVCC] + FLAGS[:2] + ["-shared", "-o", LIB] + [obj_of(s) for s in sources()])
    with open(STAMP, "w") as fh:
        fh.write(source_hash())
    return LIB


if __name__ == "__main__":
    build(force="--force" in sys.argv, verbose="-v" in sys.argv)
    print(LIB)
