"""Multi-GPU parity on real devices (run under torchrun, one rank per GPU): every sharded entry point of
cofhe_b200/shard.py must equal the same operation computed unsharded on the rank's own GPU, bit for bit."""
import os
import random
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import torch.distributed as dist

import bench as B
from conftest_free import golden_zero
from cofhe_b200 import shard as S
from cofhe_b200.group import ClassGroupB200

rank, world, local = B.dist_env()
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
fx = B.fixture("t0")
G = ClassGroupB200(fx["Delta"], local)
G.set_scheme(fx["k"], fx["N"])
G.set_bases(fx["h"], fx["pk"], fx["bound"].bit_length())
rng = random.Random(99)  # same stream on every rank: identical inputs everywhere
n = 1001
ms = G.ints([rng.getrandbits(fx["k"]) for _ in range(n)], G.ml)
rs = G.ints([rng.randrange(1, fx["bound"]) for _ in range(n)], G.el_r)
ct = G.encrypt(ms, rs)
ok = {}
ok["encrypt"] = torch.equal(S.encrypt(G, ms, rs), ct)
ok["encrypt_shared_r"] = torch.equal(S.encrypt(G, ms, 777), G.encrypt(ms, 777))
ct2 = G.encrypt(ms, 31337)
ok["add"] = torch.equal(S.add_ciphertexts(G, ct, ct2), G.add_ciphertexts(ct, ct2))
ok["part_decrypt"] = torch.equal(S.part_decrypt(G, ct, fx["sk"]), G.part_decrypt(ct, fx["sk"]))
ok["decrypt"] = torch.equal(S.decrypt(G, ct, fx["sk"]), ms)
nn, m, p = 13, 9, 5
sc = [rng.getrandbits(20) for _ in range(m * p)]
zero = G.cts([golden_zero("t0")])
ctm = ct[..., : nn * m].contiguous()
ok["scal_matmul"] = torch.equal(S.scal_matmul(G, ctm, sc, nn, m, p, zero), G.scal_matmul(ctm, sc, nn, m, p, zero))
G.status()
flag = torch.tensor([int(all(ok.values()))], device="cuda")
dist.all_reduce(flag, op=dist.ReduceOp.MIN)
if rank == 0:
    print("mgpu_check world=%d %s %s" % (world, "OK" if flag.item() else "FAIL", ok))
dist.destroy_process_group()
sys.exit(0 if flag.item() else 1)
