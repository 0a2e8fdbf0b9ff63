import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
from conftest import I, load_golden
from cofhe_b200.group import ClassGroupB200
t = load_golden("ref_t0.json"); p = t["params"]
G = ClassGroupB200(int(p["Delta"]))
G.set_scheme(int(p["k"]), int(p["N"]))
G.set_bases(I(p["h"]), I(t["pk"]), int(p["exponent_bound"]).bit_length())
ct = G.encrypt(list(range(1, 13)), 12345)
print("enc ok"); G.status()
s = G.add_ciphertexts(ct, ct); G.status(); print("add ok")
zero = G.cts([I(t["scal_matmul"][0]["zero"])])
mm = G.scal_matmul(ct, [7 * i + 3 for i in range(6)], 4, 3, 2, zero); G.status(); print("mm ok")
print(G.ints_out(G.decrypt(mm, int(t["sk"]))))
