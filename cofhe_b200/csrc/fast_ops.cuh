// cofhe_b200 -- per-element bodies of the batched entry points on the register-resident fast path (qf_fast.cuh).
//
// One thread owns one form.  Whatever the fast path declines (`bail`: see qf_fast.cuh) is appended to a per-context
// list and redone by a second, sparse kernel on the robust path (ops.cuh); both end in the unique reduced form, so
// the split is not observable in the results.
#pragma once
#include "ops.cuh"
#include "qf_fast.cuh"

#if defined(__CUDA_ARCH__)
#define QF_BLOCK_ALIGN() __syncthreads()
#define QF_BLOCK_ANY(p) (__syncthreads_or(p) != 0)
#else
#define QF_BLOCK_ALIGN() ((void)0)
#define QF_BLOCK_ANY(p) (p)
#endif

namespace cofhe {

// SoA buffer (shape s, n elements per limb row) -> registers.  Returns false when the element does not fit the
// fast shape (c wider than N+1 limbs).
template <int N>
QF_HD bool fr_load(FormR<N> &f, const u32 *p, int nc, size_t n, size_t idx) {
  QF_UNROLL for (int i = 0; i < N; i++) f.a[i] = p[(size_t)i * n + idx];
  QF_UNROLL for (int i = 0; i < N; i++) f.b[i] = p[(size_t)(N + i) * n + idx];
  QF_UNROLL for (int i = 0; i < N + 1; i++) f.c[i] = i < nc ? p[(size_t)(2 * N + i) * n + idx] : 0;
  u32 hi = 0;
  for (int i = N + 1; i < nc; i++) hi |= p[(size_t)(2 * N + i) * n + idx];
  return hi == 0;
}
// a and b only (c is redundant: see fr_fill_c)
template <int N>
QF_HD void fr_load_ab(FormR<N> &f, const u32 *p, size_t n, size_t idx) {
  QF_UNROLL for (int i = 0; i < N; i++) f.a[i] = p[(size_t)i * n + idx];
  QF_UNROLL for (int i = 0; i < N; i++) f.b[i] = p[(size_t)(N + i) * n + idx];
  QF_UNROLL for (int i = 0; i < N + 1; i++) f.c[i] = 0;
}
// c = (b^2 + |Delta|) / (4a) from a, b and the discriminant: square, one addition, and a 2-adic (Jebelean) exact
// division by the odd part of a -- about 2 % of a composition, against 13 of 27 limb rows not read at all.
// Returns false when c does not fit N+1 limbs (a tiny `a`, e.g. the neutral form: robust path).
template <int N, int CAP>
QF_HD bool fr_fill_c(FormR<N> &f, const Group<CAP> *G) {
  constexpr int W = N + 1, P = 2 * N;
  u32 bm[N];
  fw_cneg<N>(bm, f.b, f.b[N - 1] >> 31);
  u32 sq[P], D[P];
  fw_mul<N, N>(sq, bm, bm);
  const int nd = G->absD.n;
  QF_UNROLL for (int i = 0; i < P; i++) D[i] = i < nd ? G->absD.d[i] : 0u;
  u32 cy = fw_add<P>(sq, sq, D);
  if (cy || nd > P || fw_is_zero<N>(f.a)) return false;
  int ba = fw_bits<N>(f.a), bs = fw_bits<P>(sq);
  if (bs - ba - 1 > 32 * W - 2) return false;  // quotient bits <= bs - (ba + 2) + 1
  int ta = fw_ctz<N>(f.a), tz = ta + 2;
  u32 ao[N], t1[N];
  fw_copy<N>(t1, f.a);
  fw_shr_limbs<N>(t1, ta >> 5);
  fw_shr_bits<N>(ao, t1, ta & 31);
  u32 t2[P];
  fw_shr_limbs<P>(sq, tz >> 5);
  fw_shr_bits<P>(t2, sq, tz & 31);
  u32 num[W], den[W];
  QF_UNROLL for (int i = 0; i < W; i++) { num[i] = t2[i]; den[i] = i < N ? ao[i] : 0u; }
  fw_divexact_2adic<W>(f.c, num, den, inv32(ao[0]));
  return true;
}
template <int N>
QF_HD void fr_store(const FormR<N> &f, u32 *p, int nc, size_t n, size_t idx) {
  QF_UNROLL for (int i = 0; i < N; i++) p[(size_t)i * n + idx] = f.a[i];
  QF_UNROLL for (int i = 0; i < N; i++) p[(size_t)(N + i) * n + idx] = f.b[i];
  QF_UNROLL for (int i = 0; i < N + 1; i++) if (i < nc) p[(size_t)(2 * N + i) * n + idx] = f.c[i];
  for (int i = N + 1; i < nc; i++) p[(size_t)(2 * N + i) * n + idx] = 0;
}

// QFI::neg (reference qfi.inl:171-176) on registers: (a,-b,c) unless a == c or a == b
template <int N>
QF_HD void fr_inverse(FormR<N> &f) {
  u32 aw[N + 1];
  fw_ext<N + 1, N>(aw, f.a, false);
  bool a_eq_c = fw_cmp<N + 1>(aw, f.c) == 0;
  bool a_eq_b = fw_cmp<N>(f.a, f.b) == 0;  // b == a > 0: same bit pattern
  if (!(a_eq_c || a_eq_b)) fw_neg<N>(f.b, f.b);
}

// packed table entry: a[N] b[N] c[N+1] = 3N+1 words
template <int N>
QF_HD void fr_pack(u32 *t, const FormR<N> &f) {
  QF_UNROLL for (int i = 0; i < N; i++) { t[i] = f.a[i]; t[N + i] = f.b[i]; }
  QF_UNROLL for (int i = 0; i < N + 1; i++) t[2 * N + i] = f.c[i];
}
template <int N>
QF_HD void fr_unpack(FormR<N> &f, const u32 *t) {
  QF_UNROLL for (int i = 0; i < N; i++) { f.a[i] = t[i]; f.b[i] = t[N + i]; }
  QF_UNROLL for (int i = 0; i < N + 1; i++) f.c[i] = t[2 * N + i];
}

// ---- per-operation fallback -----------------------------------------------------------------------------------
// What fq_compose declines (about 2e-6 of random compositions: gcd(a1,a2) >= 2^32, a 31+ bit Euclid quotient, ...)
// is recomputed in place by the robust path through ONE out-of-line call, so a declined step costs its lane one slow
// composition instead of sending the whole element (a 100+ step exponentiation, a matmul output) to the robust
// kernels.  x: packed f1 in, packed result out; t: packed f2.  Returns false when the result does not fit the
// register shape or the robust path reports an error (the element then goes to the bail list).
template <int N, int CAP>
QF_HDN bool slow_compose(const Group<CAP> *G, u32 *x, const u32 *t, bool neg2, bool dupl) {
  Err e; e.code = 0;
  Form<CAP> f1, f2;
  z_load(e, f1.a, x, N, 1, false); z_load(e, f1.b, x + N, N, 1, true); z_load(e, f1.c, x + 2 * N, N + 1, 1, false);
  if (!dupl) { z_load(e, f2.a, t, N, 1, false); z_load(e, f2.b, t + N, N, 1, true); z_load(e, f2.c, t + 2 * N, N + 1, 1, false); }
  if (dupl) qf_compose(e, *G, f1, f1, f1, false, true);
  else qf_compose(e, *G, f1, f1, f2, neg2, false);
  z_store(e, f1.a, x, N, 1, false); z_store(e, f1.b, x + N, N, 1, true); z_store(e, f1.c, x + 2 * N, N + 1, 1, false);
  return e.code == 0;
}
template <int N, int CAP, bool DUPL, bool SYNC = false>
QF_HD bool hy_compose(const Group<CAP> *G, FormR<N> &r, const FormR<N> &f1, const FormR<N> &f2, bool neg2, int dbits) {
  bool ok = fq_compose<N, DUPL, SYNC>(r, f1, f2, neg2, dbits);
  if (!ok) {
    u32 xb[3 * N + 1], tb[3 * N + 1];
    fr_pack<N>(xb, f1);
    fr_pack<N>(tb, f2);
    ok = slow_compose<N, CAP>(G, xb, tb, neg2, DUPL);
    fr_unpack<N>(r, xb);
  }
  return ok;
}

// Out-of-line composition on packed forms, for the kernels that compose in a loop (exponentiation, matmul): one copy
// of the (large) composition body per (N, DUPL) in the whole library instead of one per call site, a kernel register
// budget that does not have to hold the loop state across it, and an instruction footprint the SM can cache.
// x: packed f1 in, packed result out (thread-local); t: packed f2 (any address space).
// SYNC: the composition runs with block barriers between its phases (qf_fast.cuh), so EVERY thread of the block must
// make the call; a thread with nothing to compose passes active = false, walks through on whatever x and t hold and
// leaves x untouched.
template <int N, int CAP, bool DUPL, bool SYNC = false, bool FLAT = false>
QF_HDN bool cc_compose(const Group<CAP> *G, u32 *x, const u32 *t, bool neg2, int dbits, bool active = true) {
  FormR<N> f1, f2, r;
  fr_unpack<N>(f1, x);
  if (!DUPL) fr_unpack<N>(f2, t);
  bool ok = fq_compose<N, DUPL, SYNC, FLAT>(r, f1, DUPL ? f1 : f2, neg2, dbits);
  if (!active) return true;
  if (ok) fr_pack<N>(x, r);
  else {
    u32 tb[3 * N + 1];
    QF_UNROLL for (int i = 0; i < 3 * N + 1; i++) tb[i] = DUPL ? 0u : t[i];
    ok = slow_compose<N, CAP>(G, x, tb, neg2, DUPL);
  }
  return ok;
}

// w-bit digit number `win` (from the least significant end) of an exponent stored with `stride` words between limbs
QF_HD u32 exp_digit(const u32 *e, int el, size_t stride, int win, int w) {
  int bit = win * w, li = bit >> 5, sh = bit & 31;
  if (li >= el) return 0;
  u64 v = e[(size_t)li * stride];
  if (li + 1 < el) v |= (u64)e[(size_t)(li + 1) * stride] << 32;
  return (u32)(v >> sh) & ((1u << w) - 1);
}

// x <- f^|e| by fixed 2^W-ary windows (left to right).  The loop is a small state machine with exactly one
// squaring site and one composition site, so the inlined composition code exists twice in the kernel, not once
// per use.  Fixed windows keep the op sequence of a warp uniform even when every thread has its own exponent
// (threads whose leading digits are zero idle until their first non-zero digit).
// tab = per-thread scratch of (2^W - 1) packed forms f^1 .. f^(2^W-1).
// Returns false when a step fails even on the robust path, or when e == 0 (the neutral form does not fit the register shape: robust path).
template <int N, int CAP, int W, bool FLAT = false>
QF_HD bool fr_pow(const Group<CAP> *G, FormR<N> &x, const FormR<N> &f, const u32 *e, int el, size_t stride, int dbits, u32 *tab) {
  constexpr int TE = 3 * N + 1, NT = (1 << W) - 1;
  int win = (32 * el + W - 1) / W - 1;
  u32 xb[TE];  // the accumulator, packed (thread-local memory: it crosses the out-of-line composition)
  fr_pack<N>(tab, f);
  fr_pack<N>(xb, f);
  bool ok = true, started = false;
  int built = 1, sq_left = 0;
  u32 d = exp_digit(e, el, stride, win, W);
  for (;;) {
    // ---- decide: 0 nothing, 1 square x, 2 compose x with table entry `src`, 3 load x from table entry `src`
    int op, src = 0;
    if (built < NT) { op = built == 1 ? 1 : 2; }
    else if (sq_left > 0) { op = started ? 1 : 0; }
    else if (d) { op = started ? 2 : 3; src = (int)d - 1; }
    else op = 0;
    if (!ok) op = 0;  // a failed element idles: every thread of the block runs the same number of rounds
    // ---- execute.  The barrier keeps the warps of a block in the same region of the (large) composition code, so
    // that the SM's instruction cache serves them together; without it they drift apart and refetch from L2.
    // Whether a step is a squaring or a table composition is the same for the whole block (it depends on the
    // window position only); who takes part is not.
    const bool sq_phase = built < NT ? built == 1 : sq_left > 0;
    if (sq_phase) {
      if (QF_BLOCK_ANY(op == 1)) { bool k = cc_compose<N, CAP, true, true, FLAT>(G, xb, xb, false, dbits, op == 1); if (op == 1) ok = k; }
    } else {
      if (QF_BLOCK_ANY(op == 2)) { bool k = cc_compose<N, CAP, false, true, FLAT>(G, xb, tab + src * TE, false, dbits, op == 2); if (op == 2) ok = k; }
    }
    if (op == 3) {
      QF_UNROLL for (int i = 0; i < TE; i++) xb[i] = tab[src * TE + i];
      started = true;
    }
    // ---- advance
    if (built < NT) {
      if (ok) { QF_UNROLL for (int i = 0; i < TE; i++) tab[built * TE + i] = xb[i]; }
      built++;
    } else if (sq_left > 0) {
      sq_left--;
    } else {
      win--;
      if (win < 0) break;
      sq_left = W;
      d = exp_digit(e, el, stride, win, W);
    }
  }
  fr_unpack<N>(x, xb);
  return ok && started;
}

// w <= 32 bits of an exponent starting at bit `bit` (from the least significant end), limbs contiguous
QF_HD u32 exp_bits(const u32 *e, int el, int bit, int w) {
  int li = bit >> 5, sh = bit & 31;
  if (li >= el) return 0;
  u64 v = e[li];
  if (li + 1 < el) v |= (u64)e[li + 1] << 32;
  return (u32)(v >> sh) & (w >= 32 ? 0xffffffffu : ((1u << w) - 1));
}

// x <- f^|e| for an exponent that the whole launch shares (part_decrypt, decrypt, tensor x scalar): left-to-right SLIDING
// windows over a table of the odd powers f, f^3, .., f^(2^W - 1).  Every thread decodes the same exponent, so the op
// sequence is uniform by construction and the windows may start wherever the next set bit is: 2^(W-1) - 1 compositions
// and one squaring build the table (fixed windows: 2^W - 3 and one), and a window costs one composition per W + 1 bits on
// average instead of one per W.  For a 105-bit share and W = 4: 104 + 8 + ~21 operations against 104 + 14 + 26.
// The same two call sites as fr_pow (one squaring, one composition), the same barrier discipline.
// tab = per-thread scratch of 2^(W-1) + 1 packed forms (the odd powers, then f^2 while the table is built).
// Returns false when a step fails even on the robust path, or when e == 0 (robust path, as fr_pow).
template <int N, int CAP, int W, bool FLAT = false>
QF_HD bool fr_pow_shared(const Group<CAP> *G, FormR<N> &x, const FormR<N> &f, const u32 *e, int el, int dbits, u32 *tab) {
  constexpr int TE = 3 * N + 1, NT = 1 << (W - 1);
  int top = -1;
  for (int i = 0; i < el; i++) if (e[i]) top = 32 * i + 31 - clz32(e[i]);
  u32 xb[TE];
  fr_pack<N>(tab, f);
  fr_pack<N>(xb, f);
  bool ok = true, started = false, have2 = false;
  // the table only as far as the exponent can use it (e < 2^W: its odd part is the only window)
  int nt = NT;
  if (top >= 0 && top < W) { u32 v = e[0] >> ctz32(e[0]); nt = (int)((v + 1) >> 1); }
  int built = 1, bit = top, sq_left = 0, pend = -1;
  while (top >= 0) {
    // ---- decide: 0 nothing, 1 square x, 2 compose x with table slot `src`, 3 load x from table slot `src`
    int op = 0, src = 0;
    if (built < nt) {
      if (!have2) op = 1;            // x = f -> f^2
      else { op = 2; src = NT; }     // x = f^(2 built - 1) o f^2
    } else {
      if (sq_left == 0 && pend < 0) {
        if (bit < 0) break;
        if (!exp_bits(e, el, bit, 1)) { sq_left = 1; bit--; }
        else {
          int l = bit + 1 < W ? bit + 1 : W;
          u32 v = exp_bits(e, el, bit - l + 1, l);  // bits bit .. bit-l+1, the top one set
          int tz = ctz32(v);
          v >>= tz; l -= tz;                         // the window ends at its lowest set bit
          sq_left = l; pend = (int)(v >> 1); bit -= l;
        }
      }
      if (sq_left > 0) op = started ? 1 : 0;         // before the first window there is nothing to square
      else { op = started ? 2 : 3; src = pend; }
    }
    const bool sq_phase = built < nt ? !have2 : sq_left > 0;
    if (!ok) op = 0;  // a failed element idles: every thread of the block runs the same number of rounds
    if (sq_phase) {
      if (QF_BLOCK_ANY(op == 1)) { bool k = cc_compose<N, CAP, true, true, FLAT>(G, xb, xb, false, dbits, op == 1); if (op == 1) ok = k; }
    } else {
      if (QF_BLOCK_ANY(op == 2)) { bool k = cc_compose<N, CAP, false, true, FLAT>(G, xb, tab + src * TE, false, dbits, op == 2); if (op == 2) ok = k; }
    }
    if (op == 3) { QF_UNROLL for (int i = 0; i < TE; i++) xb[i] = tab[src * TE + i]; }
    // ---- advance
    if (built < nt) {
      if (!have2) {
        if (ok) { QF_UNROLL for (int i = 0; i < TE; i++) { tab[NT * TE + i] = xb[i]; xb[i] = tab[i]; } }  // f^2 kept, x back to f
        have2 = true;
      } else {
        if (ok) { QF_UNROLL for (int i = 0; i < TE; i++) tab[built * TE + i] = xb[i]; }
        built++;
      }
    } else if (sq_left > 0) {
      sq_left--;
    } else {
      pend = -1;
      started = true;
    }
  }
  fr_unpack<N>(x, xb);
  return ok && started;
}

}  // namespace cofhe
