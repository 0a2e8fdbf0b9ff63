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

// QFI::neg (reference thirdparty/bicycl/src/bicycl/qfi.inl:171-176): (a, -b, c) unless a == c or a == b
template <int T, int L>
CG_FN void cq_inverse(const Grp<T> &g, FormC<L> &f, bool on) {
  u32 nb[L];
  g_neg<T, L>(g, nb, f.b);
  int ac = g_cmp<T, L>(g, f.a, f.c);
  int ab = g_cmp<T, L>(g, f.a, f.b);  // b >= 0 here whenever it can equal a (a negative b has its top bits set)
  bool bneg = g_sign<T, L>(g, f.b);
  bool keep = ac == 0 || (!bneg && ab == 0);
  g_copy_if<T, L>(f.b, nb, on && !keep);
}

// Exponentiation, one lane group per element, fixed 2^4-ary windows left to right (the op sequence is the same for every
// element, so the groups of a warp stay in lock step whatever their exponents are): table T[d] = f^d, d = 1..15, in global
// scratch (one SoA batch per entry), then per window four squarings and one composition with T[digit].  ONE inlined
// squaring and ONE inlined composition serve the table build and the main loop (state machine): the two bodies are the
// bulk of the kernel's code.  A step that declines kills the element (robust redo of the whole exponentiation).
constexpr int kCoopPowW = 4;
// the per-group body: in / out / mytab are the element batches of this form block; returns false when the element has to be
// redone by the robust path (nothing stored then)
template <int T, int L, int N>
CG_FN bool cq_pow(const Grp<T> &g, int dbits, int nc, size_t n, size_t idx, bool have, const u32 *in, u32 *out, u32 *mytab, const u32 *exp, int el,
                  int per_elem, const unsigned char *expneg, int shared_neg) {
  constexpr int W = kCoopPowW, NT = (1 << W) - 1;
  const size_t fl = (size_t)(2 * N + nc);
  FormC<L> acc, opnd, r;
  bool alive = cq_load<T, L, N>(g, acc, in, nc, n, idx) && have;
  // ---- table: T[1] = f, T[d] = T[d-1] o f
  cq_store<T, L, N>(g, acc, mytab, nc, n, idx, true);
  g_sync<T>(g);
  const int nwin = (32 * el + W - 1) / W;
  bool started = false;  // acc holds a partial power (main loop)
  // steps: 0 .. NT-2 table build (composition with T[1]); then per window W squarings and one composition
  const int nsteps = (NT - 1) + nwin * (W + 1);
  QF_NOUNROLL for (int s = 0; s < nsteps; s++) {
    const bool build = s < NT - 1;
    const int ws = build ? 0 : (s - (NT - 1)), wi = nwin - 1 - ws / (W + 1), ph = ws % (W + 1);
    const bool is_comp = build ? s > 0 : ph == W;  // T[2] = f^2 is a squaring (a form is not composed with itself)
    u32 digit = 0;
    if (!build) {
      int bit = wi * W;
      u32 limb = per_elem ? exp[(size_t)(bit >> 5) * n + idx] : exp[bit >> 5];
      digit = (limb >> (bit & 31)) & (u32)NT;
    }
    if (is_comp) {
      const bool need = build || digit != 0;
      // windows above the exponent's top digit: nothing to do anywhere in the warp
      if (!w_any<T>(g, alive && need)) continue;
      const u32 d = build ? 1u : (digit ? digit : 1u);
      cq_load<T, L, N>(g, opnd, mytab + (size_t)(d - 1) * fl * n, nc, n, idx);
      const bool first = !build && !started && digit != 0;  // the top digit: acc = T[digit]
      const bool act = alive && need && !first;
      bool ok = cq_compose<T, L, N, false>(g, r, acc, opnd, false, dbits, act);
      if (act) {
        if (ok) acc = r;
        else alive = false;
      }
      if (alive && first) { acc = opnd; started = true; }
      if (build) {
        cq_store<T, L, N>(g, acc, mytab + (size_t)(s + 1) * fl * n, nc, n, idx, alive);
        g_sync<T>(g);
      }
    } else {
      const bool act = alive && (build || started);
      if (!w_any<T>(g, act)) continue;
      bool ok = cq_compose<T, L, N, true>(g, r, acc, acc, false, dbits, act);
      if (act) {
        if (ok) acc = r;
        else alive = false;
      }
      if (build) {
        cq_store<T, L, N>(g, acc, mytab + (size_t)(s + 1) * fl * n, nc, n, idx, alive);
        g_sync<T>(g);
      }
    }
  }
  if (!started) alive = false;  // exponent 0 (the neutral form does not fit the working shape): robust path
  const bool inv = per_elem ? (expneg && expneg[idx]) : shared_neg != 0;
  cq_inverse<T, L>(g, acc, inv);
  return cq_store<T, L, N>(g, acc, out, nc, n, idx, have && alive);
}

}  // namespace cofhe
