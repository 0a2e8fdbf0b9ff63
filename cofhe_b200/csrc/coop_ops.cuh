// cofhe_b200 -- per-group bodies of the batched entry points on the cooperative path (coop_qf.cuh): SoA batches <->
// lane-blocked registers, packed table entries, the exponentiation state machine.
#pragma once
#include "coop_qf.cuh"
#include "matmul_shape.cuh"
#include "qform.cuh"

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

// One robust composition for the groups whose cooperative composition declined (`want`), in the middle of a loop of
// compositions: lane 0 of such a group collects both operands through the group's shared scratch, runs the generic
// composition (qform.cuh, dynamic-length integers in local memory) and hands the result back; the other lanes of the
// warp wait.  Slow (milliseconds at 2053 bits) but rare (~10^-6 of compositions), and the alternative -- redoing the
// whole exponentiation or matmul output on the generic path -- costs a thousand of these.  Every lane of the warp
// makes the call (the callers branch on w_any).  false: the result does not fit the working shape either.
template <int T, int L, int N, int CAP>
CG_FN bool cq_slow_compose(const Grp<T> &g, const Group<CAP> *G, FormC<L> &r, const FormC<L> &f1, const FormC<L> &f2, bool dupl, bool want) {
  constexpr int K = T * L;
  static_assert(CgScratch<T, L>::words >= 3 * K, "the group's scratch holds one form");
  u32 *s = g.scr;
  Err e; e.code = 0;
  Form<CAP> x, y;
  g_sync<T>(g);
  QF_UNROLL for (int i = 0; i < L; i++) { s[g.lane * L + i] = f1.a[i]; s[K + g.lane * L + i] = f1.b[i]; s[2 * K + g.lane * L + i] = f1.c[i]; }
  g_sync<T>(g);
  if (want && g.lane == 0) { z_load(e, x.a, s, K, 1, false); z_load(e, x.b, s + K, K, 1, true); z_load(e, x.c, s + 2 * K, K, 1, false); }
  g_sync<T>(g);
  QF_UNROLL for (int i = 0; i < L; i++) { s[g.lane * L + i] = f2.a[i]; s[K + g.lane * L + i] = f2.b[i]; s[2 * K + g.lane * L + i] = f2.c[i]; }
  g_sync<T>(g);
  if (want && g.lane == 0) {
    if (!dupl) { z_load(e, y.a, s, K, 1, false); z_load(e, y.b, s + K, K, 1, true); z_load(e, y.c, s + 2 * K, K, 1, false); }
    if (dupl) qf_compose(e, *G, x, x, x, false, true);
    else qf_compose(e, *G, x, x, y, false, false);
    // a and |b| of a reduced form fit N limbs; c must leave the guard bits of the working width free
    if (z_bits(x.a) > 32 * N - 1 || z_bits(x.c) > 32 * K - 4) e.code = 1;
    z_store(e, x.a, s, K, 1, false); z_store(e, x.b, s + K, K, 1, true); z_store(e, x.c, s + 2 * K, K, 1, false);
    s[3 * K] = (u32)e.code;
  }
  g_sync<T>(g);
  bool ok = want && s[3 * K] == 0;
  QF_UNROLL for (int i = 0; i < L; i++) {
    r.a[i] = ok ? s[g.lane * L + i] : r.a[i];
    r.b[i] = ok ? s[K + g.lane * L + i] : r.b[i];
    r.c[i] = ok ? s[2 * K + g.lane * L + i] : r.c[i];
  }
  g_sync<T>(g);
  return ok;
}

// "does any group of the BLOCK need this step": the loops below skip a step only when no warp of the block has work in
// it, so that every warp makes the same sequence of composition calls and the compositions can run with block barriers
// between their phases (cq_compose SYNC: the warps of an SM share the instruction-cache lines of the long straight-line
// code).  The host build (one simulated warp) votes inside the warp.
#if defined(__CUDACC__)
template <int T> CG_FN bool b_any(const Grp<T> &g, bool p) { (void)g; return __syncthreads_or(p ? 1 : 0) != 0; }
constexpr bool kCoopLoopSync = true;
#else
template <int T> CG_FN bool b_any(const Grp<T> &g, bool p) { return w_any<T>(g, p); }
constexpr bool kCoopLoopSync = false;
#endif

// Exponentiation, one lane group per element, fixed 2^4-ary windows left to right (the op sequence is the same for every
// element, so the groups of a warp stay in lock step whatever their exponents are): table T[d] = f^d, d = 1..15, in global
// scratch (one SoA batch per entry), then per window four squarings and one composition with T[digit].  ONE inlined
// squaring and ONE inlined composition serve the table build and the main loop (state machine): the two bodies are the
// bulk of the kernel's code.  A step that declines kills the element (robust redo of the whole exponentiation).
constexpr int kCoopPowW = 4;
// the per-group body: in / out / mytab are the element batches of this form block; returns false when the element has to be
// redone by the robust path (nothing stored then)
template <int T, int L, int N, int CAP>
CG_FN bool cq_pow(const Grp<T> &g, const Group<CAP> *G, int dbits, int nc, size_t n, size_t idx, bool have, const u32 *in, u32 *out, u32 *mytab, const u32 *exp, int el,
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
      if (!b_any<T>(g, alive && need)) continue;
      const u32 d = build ? 1u : (digit ? digit : 1u);
      cq_load<T, L, N>(g, opnd, mytab + (size_t)(d - 1) * fl * n, nc, n, idx);
      const bool first = !build && !started && digit != 0;  // the top digit: acc = T[digit]
      const bool act = alive && need && !first;
      bool ok = cq_compose<T, L, N, false, kCoopLoopSync>(g, r, acc, opnd, false, dbits, act);
      if (w_any<T>(g, act && !ok)) ok = cq_slow_compose<T, L, N, CAP>(g, G, r, acc, opnd, false, act && !ok) || ok;
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
      if (!b_any<T>(g, act)) continue;
      bool ok = cq_compose<T, L, N, true, kCoopLoopSync>(g, r, acc, acc, false, dbits, act);
      if (w_any<T>(g, act && !ok)) ok = cq_slow_compose<T, L, N, CAP>(g, G, r, acc, acc, true, act && !ok) || ok;
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

// ---- multi-exponentiation from window tables (scalar x ciphertext matmul, fixed base) on the cooperative path -------------
//   out[i,k] = (init o) prod_{j in chunk} prod_w T[digit_w(s[j,k])][w][i*m + j]          (matmul.cuh has the register version)
// The tables are ordinary SoA form batches of ne = (2^W - 1) * nwin * n*m elements, entry (d, w, base) at element
// (d-1)*nwin*nb + w*nb + base, built by the batched squaring / composition launches themselves (kernels.cu
// coop_tables); this body only composes.  One lane group per (chunk, output); the trip counts (chunk's j range, windows)
// are the same for every group of the launch, a zero digit is skipped when no group of the warp has one to do.
CG_FN u32 cq_digit(const u32 *e, int el, size_t stride, int win, int w) {
  int bit = win * w, li = bit >> 5, sh = bit & 31;
  if (li >= el) return 0;
  u64 v = e[(size_t)li * stride];
  if (li + 1 < el) v |= (u64)e[(size_t)(li + 1) * stride] << 32;
  return (u32)(v >> sh) & ((1u << w) - 1);
}
// tab / init / out: the batches of THIS form block (init: one element or null; taken by chunk 0).  Returns false when the
// output has to be redone by the robust path (a declined composition, or an empty product: the neutral form does not fit
// the working shape); a valid placeholder form is stored then, so that the launches that combine the chunks see forms.
template <int T, int L, int N, int CAP>
CG_FN bool cq_tabprod(const Grp<T> &g, const Group<CAP> *G, int dbits, int nc, const MMShape &s, unsigned chunk, size_t o, bool have, const u32 *scal,
                      const u32 *tab, size_t ne, const u32 *init, u32 *out, size_t no) {
  const size_t nb = (size_t)s.n * s.m, nbw = nb * (size_t)s.nwin, sstride = (size_t)s.m * s.p;
  const unsigned i = (unsigned)(o / s.p), k = (unsigned)(o % s.p);
  const unsigned j0 = chunk * s.jper, j1 = j0 + s.jper < s.m ? j0 + s.jper : s.m;
  FormC<L> acc, opnd, r;
  bool alive = have, started = false;
  cq_load<T, L, N>(g, acc, tab, nc, ne, (size_t)i * s.m + j0);  // T[1][0][i*m + j0]: the placeholder, a valid form
  if (init && chunk == 0) {
    bool iok = cq_load<T, L, N>(g, opnd, init, nc, 1, 0);
    if (iok) acc = opnd;
    alive = alive && iok;
    started = true;
  }
  QF_NOUNROLL for (unsigned j = j0; j < j1; j++) {
    const u32 *e = scal + (size_t)j * s.p + k;
    QF_NOUNROLL for (int w = 0; w < s.nwin; w++) {
      const u32 d = cq_digit(e, s.el, sstride, w, s.W);
      const bool need = alive && d != 0;
      if (!b_any<T>(g, need)) continue;
      const size_t ent = (size_t)((d ? d : 1u) - 1) * nbw + (size_t)w * nb + ((size_t)i * s.m + j);
      const bool lok = cq_load<T, L, N>(g, opnd, tab, nc, ne, ent);
      if (need && !lok) alive = false;
      const bool first = need && alive && !started;
      const bool act = need && alive && started;
      bool ok = cq_compose<T, L, N, false, kCoopLoopSync>(g, r, acc, opnd, false, dbits, act);
      if (w_any<T>(g, act && !ok)) ok = cq_slow_compose<T, L, N, CAP>(g, G, r, acc, opnd, false, act && !ok) || ok;
      if (act) {
        if (ok) acc = r;
        else alive = false;
      }
      if (first) { acc = opnd; started = true; }
    }
  }
  const bool good = alive && started;
  if (!good) cq_load<T, L, N>(g, acc, tab, nc, ne, (size_t)i * s.m + j0);
  const bool stored = cq_store<T, L, N>(g, acc, out, nc, no, o, have);
  return good && stored;
}

}  // namespace cofhe
