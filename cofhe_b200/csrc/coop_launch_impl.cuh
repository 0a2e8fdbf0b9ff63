// cofhe_b200 -- kernels of the cooperative path and the bodies of CoopKernels<T, L, N, CAP> (coop_launch.cuh).
// Included only by the translation units coop_n<N>.cu.
#pragma once
#include "coop_launch.cuh"
#include "coop_ops.cuh"
#include "ops.cuh"


namespace cofhe {

template <int T>
__device__ __forceinline__ Grp<T> make_group(u32 *scr_block, int words) {
  Grp<T> g;
  g.lane = threadIdx.x % T;
  g.scr = scr_block + (threadIdx.x / T) * words;
  g.base = (threadIdx.x & 31) & ~(T - 1);
  g.all = __activemask();  // every lane is here: 0xffffffff
  return g;
}

// one lane group per form operation, the 32/T groups of a warp in lock step (coop.cuh); grid.y selects the form block
// inside a ciphertext batch (0: c1, 1: c2).  Groups past the end of the batch ride along on the last element.
template <int T, int L, int N, int CAP, bool DUPL, int MB, bool SYNC = false, int BLK = kCoopBlock>
__global__ void __launch_bounds__(BLK, MB) k_coop_compose(int dbits, int nc, size_t n, const u32 *in1, const u32 *in2, u32 *out,
                                                                             int neg2, u32 *bail, int *nbail) {
  constexpr int GPB = BLK / T, SW = CgScratch<T, L>::words;
  gs_pick(dbits, blockIdx.y);  // compact variant: form block 0 (c1) lives in Cl(DeltaK)
  __shared__ u32 scr[GPB * SW];
  Grp<T> g = make_group<T>(scr, SW);
  size_t idx = (size_t)blockIdx.x * GPB + threadIdx.x / T;
  const bool have = idx < n;
  if (!have) idx = n - 1;
  size_t off = (size_t)blockIdx.y * (2 * N + nc) * n;
  FormC<L> f1, f2, r;
  bool ok = cq_load<T, L, N>(g, f1, in1 + off, nc, n, idx);
  if (!DUPL) ok = cq_load<T, L, N>(g, f2, in2 + off, nc, n, idx) && ok;
  ok = cq_compose<T, L, N, DUPL, SYNC>(g, r, f1, DUPL ? f1 : f2, (neg2 & 1) != 0, dbits, have && ok);
  ok = cq_store<T, L, N>(g, r, out + off, nc, n, idx, have && ok);
  if (have && !ok && g.lane == 0) bail[atomicAdd(nbail, 1)] = (u32)(blockIdx.y * n + idx);
}

template <int T, int L, int N, int CAP, int MB, int BLK = kCoopBlock>
__global__ void __launch_bounds__(BLK, MB) k_coop_pow(const Group<CAP> *G, int dbits, int nc, size_t n, const u32 *in, u32 *out, const u32 *exp, int el,
                                                           int per_elem, const unsigned char *expneg, int shared_neg, u32 *tab, u32 *bail,
                                                           int *nbail) {
  constexpr int GPB = BLK / T, SW = CgScratch<T, L>::words, NT = (1 << kCoopPowW) - 1;
  if (gs_pick(dbits, blockIdx.y)) G += 1;
  __shared__ u32 scr[GPB * SW];
  Grp<T> g = make_group<T>(scr, SW);
  size_t idx = (size_t)blockIdx.x * GPB + threadIdx.x / T;
  const bool have = idx < n;
  if (!have) idx = n - 1;
  const size_t fl = (size_t)(2 * N + nc), off = (size_t)blockIdx.y * fl * n;
  bool ok = cq_pow<T, L, N, CAP>(g, G, dbits, nc, n, idx, have, in + off, out + off, tab + (size_t)blockIdx.y * NT * fl * n, exp, el, per_elem, expneg,
                            shared_neg);
  if (have && !ok && g.lane == 0) bail[atomicAdd(nbail, 1)] = (u32)(blockIdx.y * n + idx);
}

template <int T, int L, int N, int CAP, int MB = 3, int BLK = kCoopBlock>
__global__ void __launch_bounds__(BLK, MB) k_coop_tabprod(const Group<CAP> *G, int dbits, int nc, MMShape s, const u32 *scal, const u32 *tab, size_t ne,
                                                               const u32 *init, u32 *part, u32 *bail, int *nbail) {
  constexpr int GPB = BLK / T, SW = CgScratch<T, L>::words;
  if (gs_pick(dbits, blockIdx.y)) G += 1;
  __shared__ u32 scr[GPB * SW];
  Grp<T> g = make_group<T>(scr, SW);
  const size_t no = (size_t)s.n * s.p, fl = (size_t)(2 * N + nc);
  size_t idx = (size_t)blockIdx.x * GPB + threadIdx.x / T;
  const bool have = idx < no;
  if (!have) idx = no - 1;
  const unsigned y = blockIdx.y, c = blockIdx.z;
  bool ok = cq_tabprod<T, L, N, CAP>(g, G, dbits, nc, s, c, idx, have, scal, tab + (size_t)y * fl * ne, ne, init ? init + (size_t)y * fl : nullptr,
                                part + ((size_t)c * gridDim.y + y) * fl * no, no);
  if (have && !ok && g.lane == 0) bail[atomicAdd(nbail, 1)] = (u32)(y * no + idx);
}

// mb: resident blocks per SM the kernel is compiled for (3: up to 168 registers, 4: 128, 5: 96) -- an occupancy /
// spilling trade-off chosen per width by the caller
template <int T, int L, int N, int CAP>
void CoopKernels<T, L, N, CAP>::compose(bool dupl, unsigned ny, cudaStream_t st, const void *G, int dbits, int nc, size_t n, const u32 *in1,
                                        const u32 *in2, u32 *out, int flags, u32 *bail, int *nbail, int mb) {
  (void)G;
  constexpr int GPB = kCoopBlock / T;
  dim3 grid((unsigned)((n + GPB - 1) / GPB), ny);
#define COOP_LAUNCH(MB)                                                                                                              \
  do {                                                                                                                               \
    if (dupl) k_coop_compose<T, L, N, CAP, true, MB><<<grid, kCoopBlock, 0, st>>>(dbits, nc, n, in1, in1, out, 0, bail, nbail);        \
    else k_coop_compose<T, L, N, CAP, false, MB><<<grid, kCoopBlock, 0, st>>>(dbits, nc, n, in1, in2, out, flags, bail, nbail);       \
  } while (0)
  if (mb >= 24) {  // 24: as 14 with 256-thread blocks, two per SM (8 warps walk through the code together)
    constexpr int GPB2 = 256 / T;
    dim3 grid2((unsigned)((n + GPB2 - 1) / GPB2), ny);
    if (dupl) k_coop_compose<T, L, N, CAP, true, 2, true, 256><<<grid2, 256, 0, st>>>(dbits, nc, n, in1, in1, out, 0, bail, nbail);
    else k_coop_compose<T, L, N, CAP, false, 2, true, 256><<<grid2, 256, 0, st>>>(dbits, nc, n, in1, in2, out, flags, bail, nbail);
  }
  else if (mb >= 14) {  // 14: the 4-blocks-per-SM build with block barriers between the phases of a composition
    if (dupl) k_coop_compose<T, L, N, CAP, true, 4, true><<<grid, kCoopBlock, 0, st>>>(dbits, nc, n, in1, in1, out, 0, bail, nbail);
    else k_coop_compose<T, L, N, CAP, false, 4, true><<<grid, kCoopBlock, 0, st>>>(dbits, nc, n, in1, in2, out, flags, bail, nbail);
  }
  else if (mb >= 5) COOP_LAUNCH(5);
  else if (mb == 4) COOP_LAUNCH(4);
  else COOP_LAUNCH(3);
#undef COOP_LAUNCH
}

template <int T, int L, int N, int CAP>
void CoopKernels<T, L, N, CAP>::tabprod(unsigned ny, cudaStream_t st, const void *G, int dbits, int nc, MMShape s, const u32 *scal, const u32 *tab, size_t ne,
                                        const u32 *init, u32 *part, u32 *bail, int *nbail, int block) {
  const size_t no = (size_t)s.n * s.p;
  if (block == 256) {  // two blocks of 256 threads per SM: 8 warps walk through the code together
    constexpr int GPB = 256 / T;
    dim3 grid((unsigned)((no + GPB - 1) / GPB), ny, s.chunks);
    k_coop_tabprod<T, L, N, CAP, 2, 256><<<grid, 256, 0, st>>>((const Group<CAP> *)G, dbits, nc, s, scal, tab, ne, init, part, bail, nbail);
  } else {
    constexpr int GPB = kCoopBlock / T;
    dim3 grid((unsigned)((no + GPB - 1) / GPB), ny, s.chunks);
    k_coop_tabprod<T, L, N, CAP><<<grid, kCoopBlock, 0, st>>>((const Group<CAP> *)G, dbits, nc, s, scal, tab, ne, init, part, bail, nbail);
  }
}

template <int T, int L, int N, int CAP>
size_t CoopKernels<T, L, N, CAP>::pow_tab_words(unsigned ny, int nc, size_t n) {
  return (size_t)ny * ((1 << kCoopPowW) - 1) * (size_t)(2 * N + nc) * n;
}
template <int T, int L, int N, int CAP>
void CoopKernels<T, L, N, CAP>::pow(unsigned ny, cudaStream_t st, const void *G, int dbits, int nc, size_t n, const u32 *in, u32 *out, const u32 *exp, int el,
                                    int per_elem, const unsigned char *expneg, int shared_neg, u32 *tab, u32 *bail, int *nbail, int block) {
  if (block == 4) {  // 128 threads, compiled for 4 blocks per SM (128 registers)
    constexpr int GPB = kCoopBlock / T;
    dim3 grid((unsigned)((n + GPB - 1) / GPB), ny);
    k_coop_pow<T, L, N, CAP, 4><<<grid, kCoopBlock, 0, st>>>((const Group<CAP> *)G, dbits, nc, n, in, out, exp, el, per_elem, expneg, shared_neg, tab, bail, nbail);
  } else if (block == 256) {
    constexpr int GPB = 256 / T;
    dim3 grid((unsigned)((n + GPB - 1) / GPB), ny);
    k_coop_pow<T, L, N, CAP, 2, 256><<<grid, 256, 0, st>>>((const Group<CAP> *)G, dbits, nc, n, in, out, exp, el, per_elem, expneg, shared_neg, tab, bail, nbail);
  } else {
    constexpr int GPB = kCoopBlock / T;
    dim3 grid((unsigned)((n + GPB - 1) / GPB), ny);
    k_coop_pow<T, L, N, CAP, 3><<<grid, kCoopBlock, 0, st>>>((const Group<CAP> *)G, dbits, nc, n, in, out, exp, el, per_elem, expneg, shared_neg, tab, bail, nbail);
  }
}

}  // namespace cofhe
