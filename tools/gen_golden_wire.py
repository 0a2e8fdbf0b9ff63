"""Golden vectors for the binary tensor wire format: bytes produced by the UNMODIFIED reference
CPUCryptoSystem::serialize_ciphertext_tensor (include/x86_64/cpu_cryptosystem.inl:322-391) through oracle/_ref, for
ciphertexts taken from the committed tier fixtures plus the neutral form (b = 0: the zero / sign-flag case).
    python tools/gen_golden_wire.py   ->  tests/golden/ref_wire.json
Runs only where /root/reference is mounted (oracle/_ref built); the fixture travels."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import I, load_golden  # noqa: E402
from oracle import qf_oracle as O, ref as R  # noqa: E402

TIERS = {"tiny": "ref_tiny.json", "t0": "ref_t0.json", "node": "ref_node.json", "t1": "ref_t1.json"}
out = {}
for tier, fn in TIERS.items():
    t = load_golden(fn)
    p = t["params"]
    D = int(p["Delta"])
    a = t["add_tensors"]
    one = O.identity(D)
    cts = list(I(a["ct1"])) + list(I(a["ct2"]))[:1] + [(one, I(a["ct1"])[0][1])]  # 2x3: five ciphertexts + a neutral c1
    r = R.Reference(int(p["N"]).bit_length(), int(p["k"]))
    flat = [x for ct in cts for f in ct for x in f]
    data = bytes(r.call("serialize_ct_tensor", 2, 3, flat))
    back = r.call("deserialize_ct_tensor", list(data))
    assert back[:3] == [2, 2, 3] and [int(x) for x in back[3:]] == flat
    assert data == O.serialize_ciphertext_tensor([2, 3], cts)
    # part-decryption-result tensor (forms) and plaintext tensor: same layout, 3 / 1 integers per element
    forms = [ct[0] for ct in cts[:5]] + [one]
    fflat = [x for f in forms for x in f]
    pdr = bytes(r.call("serialize_pdr_tensor", 2, 3, fflat))
    back = r.call("deserialize_pdr_tensor", list(pdr))
    assert back[:3] == [2, 2, 3] and [int(x) for x in back[3:]] == fflat
    assert pdr == O.serialize_int_tensor([2, 3], fflat) and O.deserialize_int_tensor(pdr, 3) == ([2, 3], fflat)
    k = int(p["k"])
    pts = [0, 1, (1 << k) - 1, -5, (1 << k) + 12345, -(1 << (k // 2)), 255, 256, (1 << 64) - 1 if k >= 64 else 77, -1, 3, (1 << (2 * k)) - 3]
    ptb = bytes(r.call("serialize_pt_tensor", 3, 4, pts))
    back = r.call("deserialize_pt_tensor", list(ptb))
    # (the reference returns the deserialised plaintext tensor flattened: cpu_cryptosystem.inl:292-317 never reshapes it back)
    assert back[:2] == [1, 12] and [int(x) for x in back[2:]] == pts
    assert ptb == O.serialize_int_tensor([3, 4], pts) and O.deserialize_int_tensor(ptb, 1) == ([3, 4], pts)
    r.close()
    out[tier] = {"Delta": str(D), "shape": [2, 3], "cts": [[[str(x) for x in f] for f in ct] for ct in cts], "hex": data.hex(),
                 "forms": [[str(x) for x in f] for f in forms], "pdr_hex": pdr.hex(),
                 "pt_shape": [3, 4], "pts": [str(x) for x in pts], "pt_hex": ptb.hex()}
with open(os.path.join(ROOT, "tests", "golden", "ref_wire.json"), "w") as fh:
    json.dump(out, fh)
print({k: len(v["hex"]) // 2 for k, v in out.items()})
