"""Small fixture reader shared by bench.py and tools/ (no pytest dependency)."""
import json
import os

ROOT = os.path.dirname(os.path.abspath(__file__))


def golden_zero(tier):
    """The reference's `zero` ciphertext (encrypt(pk, 0) of the matmul fixture) of a tier: a valid ciphertext of 0."""
    with open(os.path.join(ROOT, "tests", "golden", "ref_%s.json" % tier)) as fh:
        t = json.load(fh)
    z = t["scal_matmul"][0]["zero"]
    return tuple(tuple(int(x) for x in f) for f in z)
