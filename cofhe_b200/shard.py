"""Sharding of the batched operations over the GPUs of one box (one process per GPU, torch.distributed).

Every operation on the path is a map over independent elements -- ciphertexts of a flattened tensor, or rows of the
ciphertext matrix of a matmul (reference include/x86_64/cpu_cryptosystem_tensor_ops.inl:396-417: res[i,:] depends only
on ct[i,:] and the scalars) -- so ranks take contiguous index ranges, constants (discriminant, window tables of h and
pk, scalar matrix, shares) are replicated, and there is exactly one collective: the all-gather of the SoA result
buffers (NCCL over NVLink on the GPUs, gloo in the CPU tests).  No exchange happens during compute.
"""
import torch
import torch.distributed as dist


def partition(n, world):
    """Contiguous, balanced ranges [(lo, hi)] * world covering range(n); the first n % world ranks get one more."""
    q, r = divmod(n, world)
    out, lo = [], 0
    for k in range(world):
        hi = lo + q + (1 if k < r else 0)
        out.append((lo, hi))
        lo = hi
    return out


def _env(group=None):
    if dist.is_available() and dist.is_initialized():
        return dist.get_rank(group), dist.get_world_size(group)
    return 0, 1


def take_last(t, lo, hi, unit=1):
    """Elements [lo, hi) of the last (element) dimension, `unit` consecutive elements per index (unit = m selects
    whole rows of an n x m matrix stored row-major).  Contiguous copy: limb rows stay coalesced on the device."""
    return t[..., lo * unit:hi * unit].contiguous()


def gather_last(local, counts, group=None):
    """All-gather of SoA buffers along the last dimension; counts[k] = elements held by rank k (ragged allowed).
    Returns the full buffer on every rank."""
    rank, world = _env(group)
    if world == 1:
        return local
    mx = max(counts)
    pad = local
    if local.shape[-1] < mx:
        pad = torch.zeros(local.shape[:-1] + (mx,), dtype=local.dtype, device=local.device)
        pad[..., :local.shape[-1]] = local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad.contiguous(), group=group)
    return torch.cat([p[..., :c] for p, c in zip(parts, counts)], dim=-1)


def map_elements(fn, inputs, n, unit_in=1, unit_out=1, replicated=(), group=None):
    """Run fn(*sliced_inputs, *replicated) on this rank's range of the n indices and gather the result.

    inputs: tensors whose last dimension holds n * unit_in elements (sliced); replicated: passed through unchanged.
    fn returns a tensor whose last dimension holds (hi - lo) * unit_out elements."""
    rank, world = _env(group)
    parts = partition(n, world)
    lo, hi = parts[rank]
    local = fn(*[take_last(t, lo, hi, unit_in) for t in inputs], *replicated)
    return gather_last(local, [(b - a) * unit_out for a, b in parts], group)


# ---- the sharded entry points (G: cofhe_b200.group.ClassGroupB200 of this rank's device) -------------------------------
def add_ciphertexts(G, ct1, ct2, group=None):
    return map_elements(G.add_ciphertexts, [ct1, ct2], ct1.shape[-1], group=group)


def encrypt(G, ms, r, group=None):
    """ms: device tensor [ml, n]; r: one int (shared, reference encrypt_tensor semantics) or device tensor [el, n]."""
    n = ms.shape[-1]
    if isinstance(r, int):
        return map_elements(lambda m: G.encrypt(m, r), [ms], n, group=group)
    return map_elements(G.encrypt, [ms, r], n, group=group)


def part_decrypt(G, ct, share, group=None):
    return map_elements(lambda c: G.part_decrypt(c, share), [ct], ct.shape[-1], group=group)


def combine(G, ct, ds, group=None):
    """ds: [t, fl, n]."""
    return map_elements(G.combine, [ct, ds], ct.shape[-1], group=group)


def decrypt(G, ct, sk, group=None):
    return map_elements(lambda c: G.decrypt(c, sk), [ct], ct.shape[-1], group=group)


def scal_matmul(G, ct, scalars, n, m, p, init, group=None):
    """Row-shard ct (n x m) and the result (n x p); the scalar matrix (m x p) and `init` are replicated."""
    rank, world = _env(group)
    parts = partition(n, world)
    lo, hi = parts[rank]
    rows = hi - lo
    if rows:
        local = G.scal_matmul(take_last(ct, lo, hi, m), scalars, rows, m, p, init)
    else:
        local = ct.new_empty(ct.shape[:-1] + (0,))
    return gather_last(local, [(b - a) * p for a, b in parts], group)
