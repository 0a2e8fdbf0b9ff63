// cofhe_b200 -- per-group bodies of the batched entry points on the cooperative path (coop_qf.cuh): SoA batches <->
// lane-blocked registers, packed table entries, the exponentiation state machine.
#pragma once
#include "coop_qf.cuh"

namespace cofhe {

// SoA batch (limb rows of n elements) -> lane-blocked registers.  a, b: N limbs (b sign-extended to K), c: nc limbs.
// Returns false when c does not fit the K-limb working width minus its guard bits.
template <int T, int L, int N>
CG_FN bool cq_load(const Grp<T> &g, FormC<L> &f, const u32 *p, int nc, size_t n, size_t idx) {
  constexpr int K = T * L;
  u32 hi = 0;
  QF_UNROLL for (int i = 0; i < L; i++) {
    int q = g.lane * L + i;
    f.a[i] = q < N ? p[(size_t)q * n + idx] : 0u;
    f.b[i] = q < N ? p[(size_t)(N + q) * n + idx] : 0u;
    f.c[i] = q < nc ? p[(size_t)(2 * N + q) * n + idx] : 0u;
  }
  // sign extension of b from limb N-1
  u32 sgn = g_bcast<T>(g, f.b[(N - 1) % L], (N - 1) / L) >> 31;
  g_fill_from<T, L, N>(g, f.b, sgn ? 0xffffffffu : 0u);
  // limbs of c beyond the working width, and the guard bits of the top limb
  for (int q = K + g.lane; q < nc; q += T) hi |= p[(size_t)(2 * N + q) * n + idx];
  if (g.lane == T - 1) hi |= f.c[L - 1] >> 28;
  return !g_any<T>(g, hi != 0);
}
// registers -> SoA batch when `on`; false when not `on` or c does not fit nc limbs (nothing is written then)
template <int T, int L, int N>
CG_FN bool cq_store(const Grp<T> &g, const FormC<L> &f, u32 *p, int nc, size_t n, size_t idx, bool on) {
  constexpr int K = T * L;
  bool fits = g_high_zero<T, L>(g, f.c, nc < K ? nc : K);
  if (!on || !fits) return false;
  QF_UNROLL for (int i = 0; i < L; i++) {
    int q = g.lane * L + i;
    if (q < N) { p[(size_t)q * n + idx] = f.a[i]; p[(size_t)(N + q) * n + idx] = f.b[i]; }
    if (q < nc) p[(size_t)(2 * N + q) * n + idx] = f.c[i];
  }
  for (int q = K + g.lane; q < nc; q += T) p[(size_t)(2 * N + q) * n + idx] = 0;
  return true;
}

}  // namespace cofhe
