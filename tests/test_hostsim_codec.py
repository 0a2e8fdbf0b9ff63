"""The float <-> plaintext codec (cofhe_b200/csrc/codec.cuh, compiled with g++ by tests/hostsim) against vectors from the
compiled reference (tests/golden/codec.json, tools/gen_golden_codec.py): make_plaintext / get_float_from_plaintext bit for
bit, including what GMP's default-precision mpf does to negative, fractional and huge values, for k = 16, 64, 128, 256."""
import json
import os

import numpy as np
import pytest

import hostsim as HS
from conftest import GOLDEN

NL = 12


def load():
    with open(os.path.join(GOLDEN, "codec.json")) as fh:
        return json.load(fh)


def to_limbs(zs, nl):
    mag = np.zeros((len(zs), nl), dtype=np.uint32)
    sg = np.zeros(len(zs), dtype=np.uint8)
    for i, z in enumerate(zs):
        sg[i] = z < 0
        a = abs(z)
        for j in range(nl):
            mag[i, j] = (a >> (32 * j)) & 0xFFFFFFFF
    return mag, sg


@pytest.mark.parametrize("k", [16, 64, 128, 256])
def test_codec_matches_reference(k):
    g = load()
    t = g["k"][str(k)]
    lib = HS.load()
    fb = np.array(g["float_bits"], dtype=np.uint32)
    mag = np.zeros((len(fb), NL), dtype=np.uint32)
    sg = np.zeros(len(fb), dtype=np.uint8)
    lib.hs_codec_make(k, len(fb), HS.P(fb), NL, HS.P(mag), HS.P(sg))
    got = [(-1 if sg[i] else 1) * sum(int(mag[i, j]) << (32 * j) for j in range(NL)) for i in range(len(fb))]
    assert got == [int(v) for v in t["made"]]
    for ints, want in ((t["made"], t["floats_of_made"]), (t["ints"], t["floats_of_ints"])):
        zs = [int(v) for v in ints]
        m, s = to_limbs(zs, NL)
        ob = np.zeros(len(zs), dtype=np.uint32)
        lib.hs_codec_get(k, len(zs), NL, HS.P(m), HS.P(s), HS.P(ob))
        assert [int(x) for x in ob] == want
