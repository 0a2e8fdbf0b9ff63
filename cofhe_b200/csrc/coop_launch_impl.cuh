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
  g.mask = (T == 32 ? 0xffffffffu : ((1u << T) - 1u)) << g.base;
  return g;
}

// one lane group per form operation; grid.y selects the form block inside a ciphertext batch (0: c1, 1: c2)
template <int T, int L, int N, int CAP, bool DUPL>
__global__ void __launch_bounds__(kCoopBlock) k_coop_compose(int dbits, int nc, size_t n, const u32 *in1, const u32 *in2, u32 *out, int neg2,
                                                            u32 *bail, int *nbail) {
  constexpr int GPB = kCoopBlock / T, SW = CgScratch<T, L>::words;
  __shared__ u32 scr[GPB * SW];
  Grp<T> g = make_group<T>(scr, SW);
  size_t idx = (size_t)blockIdx.x * GPB + threadIdx.x / T;
  if (idx >= n) return;  // whole groups leave together
  size_t off = (size_t)blockIdx.y * (2 * N + nc) * n;
  FormC<L> f1, f2, r;
  bool ok = cq_load<T, L, N>(g, f1, in1 + off, nc, n, idx);
  if (!DUPL) ok = cq_load<T, L, N>(g, f2, in2 + off, nc, n, idx) && ok;
  if (ok) ok = cq_compose<T, L, N, DUPL>(g, r, f1, DUPL ? f1 : f2, (neg2 & 1) != 0, dbits);
  if (ok) ok = cq_store<T, L, N>(g, r, out + off, nc, n, idx);
  if (!ok && g.lane == 0) bail[atomicAdd(nbail, 1)] = (u32)(blockIdx.y * n + idx);
}

template <int T, int L, int N, int CAP>
void CoopKernels<T, L, N, CAP>::compose(bool dupl, unsigned ny, cudaStream_t st, const void *G, int dbits, int nc, size_t n, const u32 *in1,
                                        const u32 *in2, u32 *out, int flags, u32 *bail, int *nbail) {
  (void)G;
  constexpr int GPB = kCoopBlock / T;
  dim3 grid((unsigned)((n + GPB - 1) / GPB), ny);
  if (dupl) k_coop_compose<T, L, N, CAP, true><<<grid, kCoopBlock, 0, st>>>(dbits, nc, n, in1, in1, out, 0, bail, nbail);
  else k_coop_compose<T, L, N, CAP, false><<<grid, kCoopBlock, 0, st>>>(dbits, nc, n, in1, in2, out, flags, bail, nbail);
}

}  // namespace cofhe
