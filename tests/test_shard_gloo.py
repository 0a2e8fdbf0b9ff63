"""Host logic of the multi-GPU path (cofhe_b200/shard.py) with world_size 2 and 3 over gloo on CPU: index ranges, row
units of the matmul, ragged all-gather.  The device object is replaced by a stand-in with the same method signatures
whose arithmetic is plain tensor algebra, so that sharded == unsharded can be checked without a GPU."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from cofhe_b200 import shard as S


class StandIn:
    """Same call signatures as ClassGroupB200; elementwise integer arithmetic instead of class-group arithmetic."""

    def add_ciphertexts(self, a, b):
        return a + b

    def part_decrypt(self, ct, share):
        return ct[0] * share

    def encrypt(self, m, r):
        rr = r if isinstance(r, int) else r[:1]
        return torch.stack([m[:1] * 0 + rr, m[:1] + rr]).expand(2, 3, m.shape[-1]).contiguous()

    def combine(self, ct, ds):
        return ct[1, :1] - ds.sum(0)[:1]

    def scal_matmul(self, ct, scalars, n, m, p, init):
        s = torch.tensor(scalars, dtype=ct.dtype).reshape(m, p)
        x = ct.reshape(2, ct.shape[1], n, m)
        return (torch.einsum("ylij,jk->ylik", x, s) + init.reshape(2, -1, 1, 1)).reshape(2, ct.shape[1], n * p)


def test_partition():
    for n in (0, 1, 5, 16, 17):
        for w in (1, 2, 3, 8):
            parts = S.partition(n, w)
            assert len(parts) == w and parts[0][0] == 0 and parts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(parts, parts[1:]))
            sizes = [b - a for a, b in parts]
            assert max(sizes) - min(sizes) <= 1


def _worker(rank, world, port, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        G = StandIn()
        g = torch.Generator().manual_seed(5)
        n, m, p, fl = 7, 3, 4, 3  # ragged over 2 and 3 ranks
        ct1 = torch.randint(0, 1000, (2, fl, n), generator=g)
        ct2 = torch.randint(0, 1000, (2, fl, n), generator=g)
        assert torch.equal(S.add_ciphertexts(G, ct1, ct2), G.add_ciphertexts(ct1, ct2))
        assert torch.equal(S.part_decrypt(G, ct1, 11), G.part_decrypt(ct1, 11))
        ms = torch.randint(0, 1000, (1, n), generator=g)
        assert torch.equal(S.encrypt(G, ms, 9), G.encrypt(ms, 9))
        ds = torch.randint(0, 1000, (3, fl, n), generator=g)
        # ds is [t, fl, n]: sliced on its last dimension like every SoA buffer
        assert torch.equal(S.combine(G, ct1, ds), G.combine(ct1, ds))
        ctm = torch.randint(0, 50, (2, fl, n * m), generator=g)
        sc = [int(x) for x in torch.randint(0, 9, (m * p,), generator=g)]
        init = torch.randint(0, 50, (2, fl, 1), generator=g)
        assert torch.equal(S.scal_matmul(G, ctm, sc, n, m, p, init), G.scal_matmul(ctm, sc, n, m, p, init))
        # fewer rows than ranks: some ranks hold nothing
        ct1r = ctm[..., : 1 * m].contiguous()
        assert torch.equal(S.scal_matmul(G, ct1r, sc, 1, m, p, init), G.scal_matmul(ct1r, sc, 1, m, p, init))
        q.put((rank, "ok"))
    except Exception as e:  # noqa: BLE001
        q.put((rank, "FAIL %r" % (e,)))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_sharded_equals_unsharded(world):
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_worker, args=(r, world, port, q)) for r in range(world)]
    for p in procs:
        p.start()
    res = [q.get(timeout=120) for _ in procs]
    for p in procs:
        p.join(30)
    assert sorted(res) == [(r, "ok") for r in range(world)], res
