"""Deterministic inputs of the full-size parity cases (shared by tools/gen_golden_big.py, tests/ and bench.py): the
configurations bench.py times -- matmul at p = 256, encryption / threshold decryption at 2^16 -- regenerated from
fixed seeds so that only reference outputs (or their hashes) are committed under tests/golden/."""
import hashlib
import os
import random

import numpy as np

MM_SHAPE = (2, 256, 256)      # two ct rows of BASELINE.json config 5
SWEEP_SHAPE = (2, 16, 256)
BIG_N = 1 << 16
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "big_t0.npz")


def mm_scalars_pattern():
    n, m, p = MM_SHAPE
    return [i + 1 for i in range(m * p)]          # benchmarks/local.cpp:164-167


def mm_scalars_u128():
    n, m, p = MM_SHAPE
    rng = random.Random(20261105)
    return [rng.getrandbits(128) for _ in range(m * p)]


def sweep_scalars():
    n, m, p = SWEEP_SHAPE
    rng = random.Random(5150)
    out = []
    for i in range(m * p):
        bits = (40, 64, 97, 128)[i % 4]
        out.append(0 if rng.random() < 0.1 else rng.getrandbits(bits))
    return out


def enc_inputs(k, bound, count=BIG_N):
    rng = random.Random(42)
    ms = [rng.getrandbits(k) for _ in range(count)]
    rs = [rng.randrange(1, bound) for _ in range(count)]
    return ms, rs


def sha(a):
    return hashlib.sha256(np.ascontiguousarray(a, dtype=np.uint32).tobytes()).hexdigest()


def load():
    z = np.load(GOLDEN)
    return {k: z[k] for k in z.files}


def stored_sha(z, key):
    return bytes(z[key]).decode()
