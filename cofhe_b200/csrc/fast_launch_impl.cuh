// cofhe_b200 -- kernels of the register-resident path and the bodies of FastKernels<N, CAP> (fast_launch.cuh).
// Included only by the per-limb-count translation units fast_n<N>.cu.
#pragma once
#include "fast_launch.cuh"
#include "fast_ops.cuh"
#include "matmul.cuh"

namespace cofhe {

// grid.y selects the form block inside a ciphertext batch (0: c1, 1: c2); plain form batches use grid.y == 1
template <int N, int CAP, bool DUPL, int BLK = kBlock, bool FLAT = false>
__global__ void __launch_bounds__(BLK, 512 / BLK) k_fast_compose(const Group<CAP> *G, int dbits, int nc, size_t n, const u32 *in1, const u32 *in2, u32 *out,
                                                         int neg2, u32 *bail, int *nbail) {
  if (gs_pick(dbits, blockIdx.y)) G += 1;  // compact variant: form block 0 (c1) lives in Cl(DeltaK)
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  // threads past the end stay in the kernel (the composition contains block barriers); they redo the last element
  bool live = idx < n;
  if (!live) idx = n - 1;
  size_t off = (size_t)blockIdx.y * (2 * N + nc) * n;
  FormR<N> f1, f2, r;
  bool ok;
  if (neg2 & 2) {
    // operands without their c rows (host-pointer entry: the kernel reads pinned host memory directly and c is
    // redundant): c of the operand the composition needs is recomputed from (a, b, Delta)
    fr_load_ab<N>(f1, in1 + off, n, idx);
    if (!DUPL) fr_load_ab<N>(f2, in2 + off, n, idx);
    bool second = !DUPL && !fw_lt<N>(f1.a, f2.a);  // fq_compose reads the c of the form with the smaller a
    ok = fr_fill_c<N, CAP>(second ? f2 : f1, G);
  } else {
    ok = fr_load<N>(f1, in1 + off, nc, n, idx);
    if (!DUPL) ok = fr_load<N>(f2, in2 + off, nc, n, idx) && ok;
  }
  // an element that does not fit the register shape still walks through the composition (on its truncated limbs: every
  // loop in there is bounded) so that the block's barriers see all threads; its result is discarded
  bool cok = fq_compose<N, DUPL, true, FLAT>(r, f1, DUPL ? f1 : f2, (neg2 & 1) != 0, dbits);
  if (!live) return;
  if (ok && !cok && !(neg2 & 2)) {  // declined by the register path: one robust composition for this lane
    u32 xb[3 * N + 1], tb[3 * N + 1];
    fr_pack<N>(xb, f1);
    fr_pack<N>(tb, DUPL ? f1 : f2);
    cok = slow_compose<N, CAP>(G, xb, tb, (neg2 & 1) != 0, DUPL);
    fr_unpack<N>(r, xb);
  }
  if (ok && cok) fr_store<N>(r, out + off, nc, n, idx);
  else bail[atomicAdd(nbail, 1)] = (u32)(blockIdx.y * n + idx);
}
template <int N, int CAP, bool FLAT = false>
__global__ void __launch_bounds__(kPowBlock) k_fast_pow(const Group<CAP> *G, int dbits, int nc, size_t n, const u32 *in, u32 *out,
                                                        const u32 *exp, int el, int per_elem, const unsigned char *expneg,
                                                        int shared_neg, u32 *bail, int *nbail) {
  constexpr int W = 4;
  if (gs_pick(dbits, blockIdx.y)) G += 1;
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  // threads past the end stay in the kernel (fr_pow contains block barriers); they redo the last element and store nothing
  bool live = idx < n;
  if (!live) idx = n - 1;
  size_t off = (size_t)blockIdx.y * (2 * N + nc) * n;
  u32 tab[((1 << W) - 1) * (3 * N + 1)];
  FormR<N> f, r;
  bool ok = fr_load<N>(f, in + off, nc, n, idx);
  bool pok = per_elem ? fr_pow<N, CAP, W, FLAT>(G, r, f, exp + idx, el, n, dbits, tab) : fr_pow_shared<N, CAP, W, FLAT>(G, r, f, exp, el, dbits, tab);
  ok = ok && pok;
  if (live) {
    if (ok) {
      if (per_elem ? (expneg && expneg[idx]) : shared_neg != 0) fr_inverse<N>(r);
      fr_store<N>(r, out + off, nc, n, idx);
    } else bail[atomicAdd(nbail, 1)] = (u32)(blockIdx.y * n + idx);
  }
}

template <int FN, int FCAP>
void FastKernels<FN, FCAP>::compose(bool dupl, dim3 grid, cudaStream_t st, const void *G, int dbits, int nc, size_t n, const u32 *in1,
                                    const u32 *in2, u32 *out, int flags, u32 *bail, int *nbail) {
  const Group<FCAP> *g = (const Group<FCAP> *)G;
  // grid.x arrives sized for kBlock threads; `wide` != 0 asks for blocks of that many threads (256 / 512: 8 / 16 warps kept in
  // the same region of the code by the phase barriers) -- for batches that fill the machine several times over
  const int wide = (flags >> 8) & 0xfff;
  const bool flat = ((flags >> 20) & 1) != 0;  // experimental: branch-free scalar Euclid rounds (option "fast_flat")
  flags &= 0xff;
  if (flat) {
    if (dupl) k_fast_compose<FN, FCAP, true, kBlock, true><<<grid, kBlock, 0, st>>>(g, dbits, nc, n, in1, in1, out, flags & ~1, bail, nbail);
    else k_fast_compose<FN, FCAP, false, kBlock, true><<<grid, kBlock, 0, st>>>(g, dbits, nc, n, in1, in2, out, flags, bail, nbail);
    return;
  }
  if (wide == 256 || wide == 512) {
    dim3 g2((unsigned)((n + wide - 1) / wide), grid.y);
    if (wide == 256) {
      if (dupl) k_fast_compose<FN, FCAP, true, 256><<<g2, 256, 0, st>>>(g, dbits, nc, n, in1, in1, out, flags & ~1, bail, nbail);
      else k_fast_compose<FN, FCAP, false, 256><<<g2, 256, 0, st>>>(g, dbits, nc, n, in1, in2, out, flags, bail, nbail);
    } else {
      if (dupl) k_fast_compose<FN, FCAP, true, 512><<<g2, 512, 0, st>>>(g, dbits, nc, n, in1, in1, out, flags & ~1, bail, nbail);
      else k_fast_compose<FN, FCAP, false, 512><<<g2, 512, 0, st>>>(g, dbits, nc, n, in1, in2, out, flags, bail, nbail);
    }
    return;
  }
  if (dupl) k_fast_compose<FN, FCAP, true><<<grid, kBlock, 0, st>>>(g, dbits, nc, n, in1, in1, out, flags & ~1, bail, nbail);
  else k_fast_compose<FN, FCAP, false><<<grid, kBlock, 0, st>>>(g, dbits, nc, n, in1, in2, out, flags, bail, nbail);
}
template <int FN, int FCAP>
void FastKernels<FN, FCAP>::pow(dim3 grid, int block, cudaStream_t st, const void *G, int dbits, int nc, size_t n, const u32 *in, u32 *out,
                                const u32 *exp, int el, int per_elem, const unsigned char *expneg, int shared_neg, u32 *bail,
                                int *nbail) {
  // (the branch-free scalar Euclid rounds of the composition kernel do not pay here: measured +2 % on 8 192 exponentiations)
  k_fast_pow<FN, FCAP><<<grid, block, 0, st>>>((const Group<FCAP> *)G, dbits, nc, n, in, out, exp, el, per_elem, expneg, shared_neg,
                                                   bail, nbail);
}
template <int FN, int FCAP>
void FastKernels<FN, FCAP>::mm_chain(dim3 grid, int block, cudaStream_t st, const void *G, int dbits, int nc, MMShape s, const u32 *ct, u32 *tab) {
  k_mm_chain<FN, FCAP><<<grid, block, 0, st>>>((const Group<FCAP> *)G, dbits, nc, s, ct, tab);
}
template <int FN, int FCAP>
void FastKernels<FN, FCAP>::mm_fill(dim3 grid, int block, cudaStream_t st, const void *G, int dbits, MMShape s, u32 *tab) {
  k_mm_fill<FN, FCAP><<<grid, block, 0, st>>>((const Group<FCAP> *)G, dbits, s, tab);
}
template <int FN, int FCAP>
void FastKernels<FN, FCAP>::mm_accum(dim3 grid, int block, cudaStream_t st, const void *G, int dbits, MMShape s, const u32 *scal, const u32 *tab,
                                     u32 *part) {
  k_mm_accum<FN, FCAP><<<grid, block, 0, st>>>((const Group<FCAP> *)G, dbits, s, scal, tab, part);
}
template <int FN, int FCAP>
void FastKernels<FN, FCAP>::mm_final(dim3 grid, int block, cudaStream_t st, const void *G, int dbits, int nc, MMShape s, const u32 *init,
                                     const u32 *part, u32 *out, u32 *bail, int *nbail) {
  k_mm_final<FN, FCAP><<<grid, block, 0, st>>>((const Group<FCAP> *)G, dbits, nc, s, init, part, out, bail, nbail);
}

}  // namespace cofhe
