"""Golden vectors of the float <-> plaintext codec from the compiled reference (oracle/_ref): make_plaintext and
get_float_from_plaintext for k in {16, 64, 128, 256} on fixed float bit patterns / integers.
    python tools/gen_golden_codec.py  ->  tests/golden/codec.json   (runs where /root/reference is mounted)"""
import json, os, random, struct, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle import ref as R  # noqa: E402

def bits(x):
    return struct.unpack("<I", struct.pack("<f", x))[0]

rng = random.Random(20261105)
xs = [0.0, -0.0, 1.0, -1.0, 1.5, -1.5, -0.5, -3.9, 12345.678, -12345.678, 1e10, -1e10, 1e20, -1e20, 3e38, -3e38, -2.0 ** -64, -2.0 ** -63,
      -2.0 ** -65, 2.0 ** 64, -2.0 ** 64, -2.0 ** 63, 2.0 ** 127, -2.0 ** 127, 1e-40, -1e-40]
fb = [bits(x) for x in xs]
for _ in range(400):
    fb.append((rng.getrandbits(1) << 31) | (rng.randrange(1, 255) << 23) | rng.getrandbits(23))
for e in range(1, 255, 3):
    for s in (0, 1):
        fb += [(s << 31) | (e << 23), (s << 31) | (e << 23) | 0x7fffff]
out = {"float_bits": fb, "k": {}}
for nb, k in ((64, 16), (128, 64), (128, 128), (128, 256)):
    r = R.Reference(nb, k)
    made = r.call("make_plaintext_bits", fb)
    zs = [rng.getrandbits(k) for _ in range(300)] + [rng.getrandbits(rng.randrange(1, k + 70)) for _ in range(150)] + \
         [(1 << k) - 1, 1 << (k - 1), (1 << (k - 1)) - 1, 1 << k, 0, 1, -5, -(1 << (k - 1))]
    out["k"][str(k)] = {"made": [str(v) for v in made], "ints": [str(z) for z in zs], "floats_of_made": r.call("get_float_bits", made),
                        "floats_of_ints": r.call("get_float_bits", zs)}
    r.close()
with open(os.path.join(ROOT, "tests", "golden", "codec.json"), "w") as fh:
    json.dump(out, fh)
print(len(fb), "floats x", len(out["k"]), "values of k")
