// cofhe_b200 -- NUCOMP / NUDUPL with partial reduction and canonical reduction on cooperative integers (coop.cuh): one
// form operation per lane group, the groups of a warp in lock step.  The algorithm is the one of the register path
// (qf_fast.cuh: van der Poorten's 2x4 matrix view, single-precision Lehmer rounds with Jebelean's test, C and D columns
// from the orthogonality relations by 2-adic exact divisions, float-estimated canonical reduction) at ONE working width
// of K = T*L limbs, so the guards and their proofs carry over; what changes is where the limbs live and that every
// data-dependent decision is a predicate (coop.cuh, "WARP-CONVERGED").  Results are the unique reduced forms (reference
// thirdparty/bicycl/src/bicycl/qfi.inl:523-548), identical to ClassGroup::nucomp / nucompinv / nudupl (qfi.inl:1441-1471).
// Whatever falls outside the shapes handled here is declined (`bail`) and redone by the robust path.
#pragma once
#include "coop.cuh"

namespace cofhe {

template <int L>
struct FormC {
  u32 a[L];  // > 0
  u32 b[L];  // two's complement over K limbs
  u32 c[L];  // > 0
};

// bit length of |a| for two's complement a, as fw_sbits (a safe bound)
template <int T, int L>
CG_FN int g_sbits(const Grp<T> &g, const u32 *a) {
  bool neg = g_sign<T, L>(g, a);
  u32 t[L];
  QF_UNROLL for (int i = 0; i < L; i++) t[i] = neg ? ~a[i] : a[i];
  return g_bits<T, L>(g, t) + (neg ? 1 : 0);
}
// the 64 bits of x below bit position `bb` (x >> (bb - 64), or x << (64 - bb) when bb < 64); bb is clamped to >= 1
template <int T, int L>
CG_FN u64 g_lead64(const Grp<T> &g, const u32 *x, int bb) {
  bb = bb < 1 ? 1 : bb;
  int ti = (bb - 1) >> 5, sh = 31 - ((bb - 1) & 31);
  u32 w2 = g_limb<T, L>(g, x, ti), w1 = g_limb<T, L>(g, x, ti - 1), w0 = g_limb<T, L>(g, x, ti - 2);
  if (ti < 1) w1 = 0;
  if (ti < 2) w0 = 0;
  u64 hi = ((u64)w2 << 32) | w1;
  return sh ? ((hi << sh) | (w0 >> (32 - sh))) : hi;
}

// hgcd32 (qf_fast.cuh) without branches: QF_HG_PAIRS double steps are always executed, a step that is not accepted
// switches the rest off.  The accepted steps, and so the matrix, are those of hgcd32 -- for EXACT == true cut at the same
// cap (any prefix of the Euclidean sequence is a valid round; the next round continues).  On the cooperative path the
// lanes of a warp must not drift apart inside the scalar Euclid (coop.cuh, "WARP-CONVERGED"), and every lane of the
// warp does useful or at least harmless work in every slot.  `on`: this lane has a round to compute.
template <bool EXACT>
CG_FN Mat2 hgcd32_flat(u32 ah, u32 bh, u32 min_r, bool on) {
  u32 rp = ah, rc = bh;
  u32 Sp = 1, Tp = 0, Sc = 0, Tc = 1;
  const u32 lim = EXACT ? (min_r ? min_r : 1u) : (min_r > 2u ? min_r : 2u);
  const float kLow = 0.99999905f;
  float fprev = (float)rp * kLow;
  int j = 1;
  bool go = on;
#define CQ_HG_STEP(RP, SP, TP, RC, SC, TC, ODD)                                                        \
  {                                                                                                    \
    bool ok = go && RC >= lim;                                                                         \
    float fcur = (float)RC;                                                                            \
    float qf = fprev * rcp_approx(fcur);                                                               \
    ok = ok && (qf < 524288.0f);                                                                       \
    u32 q = hg_quotient(qf);                                                                           \
    u32 rn = RP - q * RC;                                                                              \
    hg_fixup(q, rn, RC);                                                                               \
    u32 Sn, Tn;                                                                                        \
    if (EXACT) {                                                                                       \
      u64 S64 = (u64)q * SC + SP, T64 = (u64)q * TC + TP;                                              \
      ok = ok && !((S64 | T64) >> 30);                                                                 \
      Sn = (u32)S64; Tn = (u32)T64;                                                                    \
    } else {                                                                                           \
      Sn = q * SC + SP; Tn = q * TC + TP;                                                              \
      ok = ODD ? (ok & (rn >= Tn) & (RC - rn >= Sn + SC)) : (ok & (rn >= Sn) & (RC - rn >= Tn + TC));  \
    }                                                                                                  \
    RP = ok ? rn : RP; SP = ok ? Sn : SP; TP = ok ? Tn : TP;                                           \
    fprev = fcur * kLow; /* after a refused step every later one is off: no select needed */           \
    j += ok ? 1 : 0;                                                                                   \
    go = ok;                                                                                           \
  }
  QF_UNROLL for (int it = 0; it < QF_HG_PAIRS; it++) {
    CQ_HG_STEP(rp, Sp, Tp, rc, Sc, Tc, true)   // index j odd -> even
    CQ_HG_STEP(rc, Sc, Tc, rp, Sp, Tp, false)  // index j even -> odd
  }
#undef CQ_HG_STEP
  const bool swapped = ((j - 1) & 1) != 0;  // after an odd number of accepted steps the sets have swapped roles
  Mat2 m;
  m.S0 = swapped ? Sc : Sp; m.T0 = swapped ? Tc : Tp;
  m.S1 = swapped ? Sp : Sc; m.T1 = swapped ? Tp : Tc;
  m.j = j;
  return m;
}

// Half-extended Euclid on cooperative magnitudes, the contract of fw_hxgcd (qf_fast.cuh): in (R0 >= R1, C0, C1), runs
// until the first remainder with bits <= stop_bits, returns the number of Euclidean steps.  The operands are kept
// left-justified by whole limbs, so the leading digits always sit in the top lane's top registers.  `act`: this group
// takes part (the others ride along on identity steps).
template <int T, int L>
CG_FN int g_hxgcd(const Grp<T> &g, u32 *R0, u32 *R1, u32 *C0, u32 *C1, int stop_bits, bool act, bool &bail) {
  constexpr int K = T * L;
  static_assert(L >= 3, "the leading digits are read from one lane");
  bool done = !act;
  int bits0 = g_bits<T, L>(g, R0);
  if (!done && bits0 == 0) { bail = true; done = true; }
  int lsh = done ? 0 : K - (bits0 + 31) / 32;
  g_shl_limbs<T, L>(g, R0, lsh);
  g_shl_limbs<T, L>(g, R1, lsh);
  int sb = stop_bits + 32 * lsh;
  int steps = 0;
  for (int guard = 0;; guard++) {
    // renormalise by whole limbs (a round removes at most ~30 bits, so the slide runs 0 or 1 times)
    u32 r0h;
    for (;;) {
      r0h = g_bcast<T>(g, R0[L - 1], T - 1);
      bool need = !done && r0h == 0 && lsh < K - 1;
      if (!w_any<T, 11>(g, need)) break;
      g_slide_up_if<T, L>(g, R0, 0, need);
      g_slide_up_if<T, L>(g, R1, 0, need);
      if (need) { lsh++; sb += 32; }
    }
    u32 r0m = g_bcast<T>(g, R0[L - 2], T - 1), r1h = g_bcast<T>(g, R1[L - 1], T - 1), r1m = g_bcast<T>(g, R1[L - 2], T - 1);
    int b1 = 0;
    {
      bool deep = !done && !r1h && !r1m;
      if (w_any<T, 12>(g, deep)) {
        int bb = g_bits<T, L>(g, R1);
        if (deep) b1 = bb;
      }
      if (r1h) b1 = 32 * K - clz32(r1h);
      else if (r1m) b1 = 32 * (K - 1) - clz32(r1m);
    }
    if (!done) {
      if (b1 <= sb) done = true;  // includes R1 == 0
      else if (guard >= 80 * K) { bail = true; done = true; }
    }
    if (!w_any<T, 13>(g, !done)) break;
    Mat2 M;
    {
      const bool exact = lsh == K - 1;
      int sh = (exact || !r0h) ? 0 : clz32(r0h);
      u32 ah = r0h, bh = r1h;
      if (sh) { ah = (ah << sh) | (r0m >> (32 - sh)); bh = (bh << sh) | (r1m >> (32 - sh)); }
      int h = 32 * (K - 1) - sh;
      u32 min_r = (sb > h && sb - h < 32) ? (1u << (sb - h)) : 0u;
      // the truncated variant in every lane (branch-free); the exact one (last rounds of a full gcd) only in the rounds
      // where some group of the warp needs it
      Mat2 Mt = hgcd32_flat<false>(ah, bh, min_r, !done && !exact);
      M = Mt;
      if (w_any<T>(g, !done && exact)) {
        Mat2 Me = hgcd32_flat<true>(ah, bh, min_r, !done && exact);
        M.S0 = exact ? Me.S0 : Mt.S0; M.T0 = exact ? Me.T0 : Mt.T0;
        M.S1 = exact ? Me.S1 : Mt.S1; M.T1 = exact ? Me.T1 : Mt.T1;
        M.j = exact ? Me.j : Mt.j;
      }
      if (done) M.j = 0;
      QF_STAT(0);
    }
    const bool full = !done && M.j == 1;
    if (w_any<T, 14>(g, full)) {
      QF_STAT(1);
      // one full-precision division step with a quotient of at most 30 bits (rare), for the groups with `full`
      int b0 = 32 * K - clz32(r0h);
      bool ok = full && b0 - b1 <= 29 && r0h != 0;
      int s2 = clz32(r0h) & 31;
      u64 A64 = ((u64)r0h << 32) | r0m, B64 = ((u64)r1h << 32) | r1m;
      u32 a0 = g_bcast<T>(g, R0[L - 3], T - 1), b0l = g_bcast<T>(g, R1[L - 3], T - 1);
      if (s2) {
        A64 = (A64 << s2) | (a0 >> (32 - s2));
        B64 = (B64 << s2) | (b0l >> (32 - s2));
      }
      u64 q64 = A64 / (B64 == ~(u64)0 ? (u64)1 : B64 + 1);  // every lane divides (no call under a predicate)
      u32 q = (ok && B64 != ~(u64)0) ? (u32)q64 : 0u;
      u32 t[L], cn[L];
      g_copy<T, L>(t, R0);
      if (g_submul1<T, L>(g, t, R1, q) != 0) ok = false;  // q underestimates
      int fix = 0;
      for (;;) {
        int cmp = g_cmp<T, L>(g, t, R1);
        bool again = ok && cmp >= 0;
        if (!w_any<T>(g, again)) break;
        g_sub_if<T, L>(g, t, t, R1, again);
        if (again) { q++; if (++fix > 8) ok = false; }
      }
      u32 top = g_mul1<T, L>(g, cn, C1, q);
      u32 cy = g_add<T, L>(g, cn, cn, C0);
      if ((top | cy) != 0) ok = false;
      if (full) {
        if (ok) {
          g_copy<T, L>(R0, R1); g_copy<T, L>(R1, t);
          g_copy<T, L>(C0, C1); g_copy<T, L>(C1, cn);
          steps++;
        } else { bail = true; done = true; }
      }
    }
    {
      // the round's 2x2 matrix on (R0, R1) and (C0, C1); groups without one apply the identity
      const bool lin = !done && !full;
      const u32 S0 = lin ? M.S0 : 1u, T0 = lin ? M.T0 : 0u, S1 = lin ? M.S1 : 0u, T1 = lin ? M.T1 : 1u;
      const bool odd = lin ? (M.j & 1) != 0 : true;
      u32 t0[L], t1[L];
      // j odd : X_{j-1} = S0*A - T0*B >= 0 , X_j = T1*B - S1*A >= 0 ; j even: signs flipped
      g_lin_sub<T, L>(g, t0, S0, R0, T0, R1, !odd);
      g_lin_sub<T, L>(g, t1, S1, R0, T1, R1, odd);
      g_copy<T, L>(R0, t0); g_copy<T, L>(R1, t1);
      u32 o0 = g_lin_add<T, L>(g, t0, S0, C0, T0, C1);
      u32 o1 = g_lin_add<T, L>(g, t1, S1, C0, T1, C1);
      g_copy<T, L>(C0, t0); g_copy<T, L>(C1, t1);
      if (lin) {
        if (o0 | o1) { bail = true; done = true; }
        else steps += M.j - 1;
      }
    }
  }
  g_shr_limbs<T, L>(g, R0, lsh);
  g_shr_limbs<T, L>(g, R1, lsh);
  return steps;
}

// canonical reduction (qfi.inl:523-548): a, c unsigned, b two's complement, all K limbs; `act`: this group takes part;
// false = declined
template <int T, int L>
CG_FN bool cq_reduce(const Grp<T> &g, u32 *a, u32 *b, u32 *c, int dbits, bool act) {
  constexpr int K = T * L;
  bool active = act, good = false;
  for (int it = 0; it < 64; it++) {
    if (!w_any<T>(g, active)) break;
    // ---- normalise: -a < b <= a
    int ba = g_bits<T, L>(g, a);
    if (active && (ba == 0 || ba > 32 * K - 3 || ba < dbits - 32 * K + 3)) active = false;
    bool bneg = g_sign<T, L>(g, b);
    u32 bm[L];
    g_cneg<T, L>(g, bm, b, bneg);
    int bb = g_bits<T, L>(g, bm);
    bool norm = active && bb >= ba;
    if (norm && bb - ba > 28) { active = false; norm = false; }
    if (w_any<T>(g, norm)) {
      u64 B64 = g_lead64<T, L>(g, bm, bb), A64 = g_lead64<T, L>(g, a, bb);
      double kd = (double)B64 / (2.0 * (double)A64 + 2.0);
      u32 k = norm ? (u32)kd : 0u;  // underestimate of floor(|b| / 2a), < 2^29
      u32 a2[L], r[L], cn[L];
      g_shl_bits<T, L>(g, a2, a, 1);
      g_copy<T, L>(r, bm);
      if (g_submul1<T, L>(g, r, a2, k) != 0 && norm) { active = false; norm = false; }
      // bring r into (-a, a] for b > 0, [-a, a) for b < 0  (r is signed from here)
      int fix = 0;
      for (;;) {
        bool sg = g_sign<T, L>(g, r);
        int cmp = g_cmp<T, L>(g, r, a);
        bool over = norm && !sg && (bneg ? cmp >= 0 : cmp > 0);
        if (!w_any<T>(g, over)) break;
        g_sub_if<T, L>(g, r, r, a2, over);
        if (over) { k++; if (++fix > 8) { active = false; norm = false; } }
      }
      // b' = sign * r ; c' = c - k * (|b| + r)/2
      u32 hs[L], nb[L];
      g_addm<T, L>(g, hs, bm, r);
      g_shr_bits<T, L>(g, hs, hs, 1);
      g_copy<T, L>(cn, c);
      if (g_submul1<T, L>(g, cn, hs, norm ? k : 0u) != 0 && norm) { active = false; norm = false; }
      g_cneg<T, L>(g, nb, r, bneg);
      g_copy_if<T, L>(c, cn, norm);
      g_copy_if<T, L>(b, nb, norm);
    }
    // ---- finished, or a rho step if a > c
    int cmp = g_cmp<T, L>(g, a, c);
    bool sgb = g_sign<T, L>(g, b);
    bool fin = active && cmp <= 0, rho = active && cmp > 0;
    u32 nb2[L];
    g_neg<T, L>(g, nb2, b);
    g_copy_if<T, L>(b, nb2, rho || (fin && cmp == 0 && sgb));
    QF_UNROLL for (int i = 0; i < L; i++) { u32 t = a[i]; a[i] = rho ? c[i] : a[i]; c[i] = rho ? t : c[i]; }
    if (fin) { good = true; active = false; }
  }
  return good;
}

// registers -> scratch slot, product through the streaming multiplier, low K limbs back to registers; nb: limbs of b
// that matter, rounded up to a multiple of L inside (b is K limbs wide, so the padding limbs are real limbs of b)
template <int T, int L>
CG_FN void cq_mul_lo(const Grp<T> &g, u32 *r, const u32 *a, const u32 *b, int nb) {
  g_sync<T>(g);
  g_to_scr<T, L>(g, scr_a<T, L>(g), a);
  g_to_scr<T, L>(g, scr_b<T, L>(g), b);
  g_mul_s<T, L>(g, scr_a<T, L>(g), scr_b<T, L>(g), (nb + L - 1) / L * L, scr_out<T, L>(g));
  g_from_scr<T, L>(g, r, scr_out<T, L>(g));
}
// r = u mod d for K-limb u; false if d == 0
template <int T, int L>
CG_FN bool cq_mod(const Grp<T> &g, u32 *r, const u32 *u, const u32 *d) {
  g_sync<T>(g);
  g_to_scr<T, L>(g, scr_out<T, L>(g), u);
  g_to_scr<T, L>(g, scr_a<T, L>(g), d);
  bool ok = g_mod_s<T, L>(g, scr_out<T, L>(g), T * L, scr_a<T, L>(g), scr_b<T, L>(g));
  g_from_scr<T, L>(g, r, scr_b<T, L>(g));
  return ok;
}
// r = (x * y) mod d, x and y unsigned K-limb values
template <int T, int L>
CG_FN bool cq_mulmod(const Grp<T> &g, u32 *r, const u32 *x, const u32 *y, const u32 *d) {
  constexpr int K = T * L;
  g_sync<T>(g);
  g_to_scr<T, L>(g, scr_a<T, L>(g), x);
  g_to_scr<T, L>(g, scr_b<T, L>(g), y);
  g_mul_s<T, L>(g, scr_a<T, L>(g), scr_b<T, L>(g), K, scr_out<T, L>(g));
  g_to_scr<T, L>(g, scr_a<T, L>(g), d);
  bool ok = g_mod_s<T, L>(g, scr_out<T, L>(g), 2 * K, scr_a<T, L>(g), scr_b<T, L>(g));
  g_from_scr<T, L>(g, r, scr_b<T, L>(g));
  return ok;
}

// r = f1 o f2 (neg2: f1 o f2^-1; DUPL: f1^2, f2 ignored).  N = limbs of a reduced a or b.  `act`: this group has an
// element.  false = declined (or no element).
// SYNC: block barriers between the phases (only where every warp of the block makes the same number of calls: the batched
// composition kernel).  The warps of an SM then walk through the long straight-line code together and share its
// instruction-cache lines instead of evicting each other's (the register path does the same: qf_fast.cuh QF_PHASE).
#if defined(__CUDACC__)
#define CQ_PHASE() do { if (SYNC) __syncthreads(); } while (0)
#else
#define CQ_PHASE() ((void)0)
#endif
template <int T, int L, int N, bool DUPL, bool SYNC = false>
CG_FN bool cq_compose(const Grp<T> &g, FormC<L> &r, const FormC<L> &f1, const FormC<L> &f2, bool neg2, int dbits, bool act) {
  constexpr int K = T * L;
  constexpr int M = (N + 1) / 2 + 1;  // half width of the post-Euclid quantities
  static_assert(K >= 2 * M && K >= N + 2, "working width too small for this limb count");
  bool bail = !act;

  // ---- operands, larger `a` first
  u32 a1[L], a2[L], b1[L], b2[L], c2[L];
  {
    bool sw = !DUPL && g_cmp<T, L>(g, f1.a, f2.a) < 0;
    QF_UNROLL for (int i = 0; i < L; i++) {
      a1[i] = sw ? f2.a[i] : f1.a[i];
      b1[i] = sw ? f2.b[i] : f1.b[i];
      a2[i] = (DUPL || sw) ? f1.a[i] : f2.a[i];
      b2[i] = (DUPL || sw) ? f1.b[i] : f2.b[i];
      c2[i] = (DUPL || sw) ? f1.c[i] : f2.c[i];
    }
    if (!DUPL) {
      g_cneg<T, L>(g, b1, b1, neg2 && sw);
      g_cneg<T, L>(g, b2, b2, neg2 && !sw);
    }
  }
  // s = (b1+b2)/2, m = (b2-b1)/2  (two's complement)
  u32 s[L], m[L];
  g_addm<T, L>(g, s, b1, b2);
  g_shr_bits<T, L>(g, s, s, 1, g_sign<T, L>(g, s) ? 0xffffffffu : 0u);
  g_subm<T, L>(g, m, b2, s);

  CG_MARK(10);
  CQ_PHASE();
  // ---- phases 1-3 share ONE inlined instance of the Euclid (its code is the largest single piece of the kernel):
  //   phase 1: F = gcd(a1, a2) with v*a2 == F (mod a1)   [DUPL: F = a, G = gcd(a, b) = . + y*|b|]
  //   phase 2: By = a1/G and Bx
  //   phase 3: partial Euclid on (By, Bx) down to about sqrt(a1/a2) * |D|^(1/4)
  u32 By[L], Bx[L], Cv[L], C0[L], C1[L];
  u32 G = 1, F = 1;
  int st1 = 0, st = 0;
  const bool bneg = g_sign<T, L>(g, b1);
  const int ba1 = g_bits<T, L>(g, a1), ba2 = g_bits<T, L>(g, a2);
  QF_NOUNROLL for (int ph = 0; ph < 2; ph++) {
    int stop = 0;
    if (ph == 0) {
      g_copy<T, L>(By, a1);
      if (!DUPL) g_copy<T, L>(Bx, a2);
      else {
        g_cneg<T, L>(g, Bx, b1, bneg);
        if (g_is_zero<T, L>(g, Bx)) bail = true;     // b == 0 (ambiguous forms): robust path
        if (g_cmp<T, L>(g, By, Bx) < 0) bail = true;  // unreduced input
      }
    } else {
      CG_MARK(20);
      // ---- phase 2: By = a1/G and Bx
      if (!DUPL) {
        bool vneg = (st1 & 1) == 0;  // sign(C_st) = (-1)^(st+1)
        bool mneg = g_sign<T, L>(g, m);
        u32 mm[L];
        g_cneg<T, L>(g, mm, m, mneg);
        u32 y0m = 0;
        bool y0neg = false, x0neg = false, x0_is_big = false;
        u32 x0big[L];
        g_zero<T, L>(x0big);
        if (w_any<T>(g, F != 1)) {
          CG_MARK(24);
          // G = gcd(F, s) = x0*F + y0*s   (groups with F == 1 ride along: G = 1, y0 = 0, x0 unused)
          bool sneg = g_sign<T, L>(g, s);
          u32 sm_[L];
          g_cneg<T, L>(g, sm_, s, sneg);
          CG_MARK(25);
          u32 srem = g_mod1<T, L>(g, sm_, F);
          CG_MARK(26);
          // exact 32-bit Euclid with one cofactor, every group in the same loop (F == 1 or srem == 0: no steps, G = F, y = 0)
          i64 y = 0;
          {
            u32 r0 = F, r1 = srem;
            i64 y0 = 0, y1 = 1;
            while (w_any<T>(g, r1 != 0)) {
              const bool on = r1 != 0;
              const u32 dv = r1 | (on ? 0u : 1u);
              u32 q = on ? r0 / dv : 0u;
              u32 r2 = r0 - q * r1;
              i64 y2 = y0 - (i64)q * y1;
              r0 = on ? r1 : r0; r1 = on ? r2 : r1; y0 = on ? y1 : y0; y1 = on ? y2 : y1;
            }
            G = r0;            // y0 * |s| == G (mod F)
            y = sneg ? -y0 : y0;
            x0_is_big = F != 1 && srem != 0;
            if (!x0_is_big) y = 0;
          }
          CG_MARK(27);
          y0neg = y < 0;
          y0m = (u32)(y0neg ? -y : y);
          // x0 = (G - y*s)/F  (exact)
          u32 ys[L], g1[L], num[L], nm2[L];
          if (g_mul1<T, L>(g, ys, sm_, y0m) != 0 && x0_is_big) bail = true;
          bool ysz = g_is_zero<T, L>(g, ys);
          bool ysneg = (y0neg != sneg) && !ysz;
          CG_MARK(28);
          g_set_u32<T, L>(g, g1, G);
          // num = |G - (ysneg ? -ys : ys)|, nneg = sign
          bool lt = g_cmp<T, L>(g, g1, ys) < 0;
          bool nneg = !ysneg && lt;
          {
            u32 yy[L];
            g_cneg<T, L>(g, yy, ys, !ysneg);  // G + ys  or  G - ys
            g_addm<T, L>(g, num, g1, yy);
            g_cneg<T, L>(g, nm2, num, nneg);
          }
          CG_MARK(29);
          g_divexact1<T, L>(g, num, nm2, F);
          CG_MARK(291);
          g_copy_if<T, L>(x0big, num, x0_is_big);
          x0neg = nneg;
        }
        {
          u32 q[L];
          g_divexact1<T, L>(g, q, a1, G);
          g_copy<T, L>(By, q);
        }
        CG_MARK(21);
        // t = (|v|*|m|) mod By
        u32 t[L];
        if (!cq_mulmod<T, L>(g, t, Cv, mm, By)) bail = true;
        bool tneg = vneg != mneg;
        if (w_any<T>(g, x0_is_big)) {
          CG_MARK(22);
          // t <- x0 * t (mod By)
          u32 xr[L], t2[L];
          if (!cq_mod<T, L>(g, xr, x0big, By) && x0_is_big) bail = true;
          if (!cq_mulmod<T, L>(g, t2, xr, t, By) && x0_is_big) bail = true;
          g_copy_if<T, L>(t, t2, x0_is_big);
          if (x0_is_big) tneg = tneg != x0neg;
        }
        if (w_any<T>(g, y0m != 0)) {
          CG_MARK(23);
          // u = y0 * (c2 mod By) (mod By)
          const bool on = y0m != 0;
          u32 cr[L], u[L], t2[L];
          if (!cq_mod<T, L>(g, cr, c2, By) && on) bail = true;
          if (g_mul1<T, L>(g, cr, cr, y0m) != 0 && on) bail = true;
          if (!cq_mod<T, L>(g, u, cr, By) && on) bail = true;
          // Bx = (tneg ? -t : t) + (y0neg ? -u : u)  (mod By): bring both to [0, By)
          g_copy<T, L>(t2, t);
          {
            u32 w[L];
            bool tz = g_is_zero<T, L>(g, t2);
            g_subm<T, L>(g, w, By, t2);
            g_copy_if<T, L>(t2, w, tneg && !tz);
            bool uz = g_is_zero<T, L>(g, u);
            g_subm<T, L>(g, w, By, u);
            g_copy_if<T, L>(u, w, y0neg && !uz);
          }
          u32 cy = g_add<T, L>(g, t2, t2, u);
          int cge = g_cmp<T, L>(g, t2, By);
          bool ge = cy || cge >= 0;
          g_sub_if<T, L>(g, t2, t2, By, ge);
          g_copy_if<T, L>(t, t2, on);
          if (on) tneg = false;
        }
        {
          u32 w[L];
          bool tz = g_is_zero<T, L>(g, t);
          g_subm<T, L>(g, w, By, t);
          g_copy_if<T, L>(t, w, tneg && !tz);
        }
        g_copy<T, L>(Bx, t);
      } else {
        // NUDUPL: s = b, m = 0, F = a.  G = gcd(a, b) = . + y*b ; Bx = y*c mod (a/G)
        G = F;
        bool yneg = ((st1 & 1) == 0) != bneg;
        {
          u32 q[L];
          g_divexact1<T, L>(g, q, a1, G);
          g_copy<T, L>(By, q);
        }
        u32 cr[L], t[L], w[L];
        if (!cq_mod<T, L>(g, cr, c2, By)) bail = true;
        if (!cq_mulmod<T, L>(g, t, Cv, cr, By)) bail = true;
        bool tz = g_is_zero<T, L>(g, t);
        g_subm<T, L>(g, w, By, t);
        g_copy_if<T, L>(t, w, yneg && !tz);
        g_copy<T, L>(Bx, t);
      }

      CG_MARK(30);
      stop = (ba1 - ba2) / 2 + dbits / 4;
      if (stop < 1) stop = 1;
      if (g_cmp<T, L>(g, By, Bx) < 0) bail = true;  // cannot happen for Bx = . mod By; keeps the Euclid's precondition
    }
    g_zero<T, L>(C0);
    g_set_u32<T, L>(g, C1, 1);
    CQ_PHASE();
    int stp = g_hxgcd<T, L>(g, By, Bx, C0, C1, stop, !bail, bail);  // (By, Bx) <- (R_{i-1}, R_i)
    if (ph == 0) {
      st1 = stp;
      if (g_bits<T, L>(g, By) > 32) bail = true;  // gcd >= 2^32: robust path
      F = g_bcast<T>(g, By[0], 0);
      if (bail || F == 0) F = 1;
      g_copy<T, L>(Cv, C0);
    } else st = stp;
  }

  CG_MARK(40);
  CQ_PHASE();
  // ---- phase 4: the C and D columns from the orthogonality relations
  // rows: x' = (Ax, Bxp) with Ax = G*C0 >= 0, Bxp = (-1)^idx R_{i-1};  y' = (Ay, Byp) with Ay = (-1)^(idx+1) G*C1, Byp = R_i
  const int idx = 1 + st;
  const bool ay_neg = (idx & 1) == 0, rx_neg = (idx & 1) != 0;
  u32 *Rx = By, *Ry = Bx;  // magnitudes
  u32 Ax[L], Ay[L];
  u32 Cx[L], Cy[L], Dx[L], Dy[L];
  {
    const int lim = 32 * M - 2;
    int bRx = g_bits<T, L>(g, Rx), bRy = g_bits<T, L>(g, Ry);
    if (bRx > lim || bRy > lim) bail = true;
    if (g_mul1<T, L>(g, Ax, C0, G) != 0) bail = true;
    if (g_mul1<T, L>(g, Ay, C1, G) != 0) bail = true;
    int bAx = g_bits<T, L>(g, Ax), bAy = g_bits<T, L>(g, Ay);
    if (bAx > lim || bAy > lim) bail = true;
    // size guards for the quotients (they must fit M limbs signed):
    //   |C| <= (a2*|B| + |m|*|A|)/a1 ,  |D| <= (|s|*|B| + c2*|A|)/a1
    int bm = g_sbits<T, L>(g, m), bs = g_sbits<T, L>(g, s), bc2 = g_bits<T, L>(g, c2);
    int mxA = bAx > bAy ? bAx : bAy, mxR = bRx > bRy ? bRx : bRy;
    int t1 = ba2 + mxR, t2 = bm + mxA, t3 = bs + mxR, t4 = bc2 + mxA;
    int mx = t1 > t2 ? t1 : t2;
    mx = mx > t3 ? mx : t3;
    mx = mx > t4 ? mx : t4;
    if (mx + 2 - (ba1 - 1) > 32 * M - 3) bail = true;
    // 2-adic exact division by a1 = 2^tz * a1o: only the low limbs of the numerators are formed
    int tz = g_ctz<T, L>(g, a1);
    if (tz > 32 * (K - M) - 1) { bail = true; tz = 0; }
    u32 a1o[L];
    g_copy<T, L>(a1o, a1);
    g_shr_limbs<T, L>(g, a1o, tz >> 5);
    g_shr_bits<T, L>(g, a1o, a1o, tz & 31);
    // quotient Q = (p*B - q*A)/a1 for (p, q) = (a2, m) -> C and (s, c2) -> D; B, A given as magnitude + sign
    auto ortho = [&](u32 *Q, const u32 *P, const u32 *Qn, const u32 *Bmag, bool Bneg, const u32 *Amag, bool Aneg) {
      u32 n1[L], n2[L];
      cq_mul_lo<T, L>(g, n1, P, Bmag, M);
      g_cneg<T, L>(g, n1, n1, Bneg);
      cq_mul_lo<T, L>(g, n2, Qn, Amag, M);
      g_cneg<T, L>(g, n2, n2, Aneg);
      g_subm<T, L>(g, n1, n1, n2);
      g_shr_limbs<T, L>(g, n1, tz >> 5);
      g_shr_bits<T, L>(g, n1, n1, tz & 31);
      g_sync<T>(g);
      g_to_scr<T, L>(g, scr_a<T, L>(g), n1);
      g_to_scr<T, L>(g, scr_b<T, L>(g), a1o);
      g_divexact2_s<T, L>(g, scr_a<T, L>(g), scr_b<T, L>(g), (M + L - 1) / L * L, scr_out<T, L>(g));
      g_from_scr<T, L>(g, Q, scr_out<T, L>(g));
      // the quotient is an M-limb two's complement number: extend it to the working width
      u32 sgn = g_bcast<T>(g, Q[(M - 1) % L], (M - 1) / L) >> 31;
      g_fill_from<T, L, M>(g, Q, sgn ? 0xffffffffu : 0u);
    };
    if (!DUPL) ortho(Cy, a2, m, Ry, false, Ay, ay_neg);
    ortho(Dy, s, c2, Ry, false, Ay, ay_neg);
    if (!DUPL) ortho(Cx, a2, m, Rx, rx_neg, Ax, false);
    ortho(Dx, s, c2, Rx, rx_neg, Ax, false);
    if (DUPL) {
      // C column equals the B column for a square
      g_copy<T, L>(Cy, Ry);
      g_cneg<T, L>(g, Cx, Rx, rx_neg);
    }
  }

  CG_MARK(50);
  CQ_PHASE();
  // ---- phase 5: the composed form (two's complement):
  //   a3 = By*Cy - Ay*Dy ; c3 = Bx*Cx - Ax*Dx ; b3 = (Ax*Dy + Ay*Dx) - (Bx*Cy + By*Cx)
  u32 ra[L], rb[L], rc[L];
  {
    bool ncx = g_sign<T, L>(g, Cx), ncy = g_sign<T, L>(g, Cy), ndx = g_sign<T, L>(g, Dx), ndy = g_sign<T, L>(g, Dy);
    g_cneg<T, L>(g, Cx, Cx, ncx); g_cneg<T, L>(g, Cy, Cy, ncy); g_cneg<T, L>(g, Dx, Dx, ndx); g_cneg<T, L>(g, Dy, Dy, ndy);
    u32 p[L];
    auto term = [&](u32 *acc, const u32 *x, const u32 *y, bool neg, bool first) {
      cq_mul_lo<T, L>(g, p, x, y, M);
      g_cneg<T, L>(g, p, p, neg);
      if (first) g_copy<T, L>(acc, p);
      else g_addm<T, L>(g, acc, acc, p);
    };
    term(ra, Ry, Cy, ncy, true);
    term(ra, Ay, Dy, !(ay_neg != ndy), false);
    term(rc, Rx, Cx, rx_neg != ncx, true);
    term(rc, Ax, Dx, !ndx, false);
    term(rb, Ax, Dy, ndy, true);
    term(rb, Ay, Dx, ay_neg != ndx, false);
    term(rb, Rx, Cy, !(rx_neg != ncy), false);
    term(rb, Ry, Cx, !ncx, false);
  }
  // must fit the working shape: 0 < a3 < 2^(32N-1), 0 < c3 < 2^(32K-2)
  {
    bool sa = g_sign<T, L>(g, ra), sc = g_sign<T, L>(g, rc);
    if (sa || sc) bail = true;
  }
  {
    int b3a = g_bits<T, L>(g, ra), b3c = g_bits<T, L>(g, rc);
    if (b3a == 0 || b3a > 32 * N - 1 || b3c == 0 || b3c > 32 * K - 3) bail = true;
  }

  CG_MARK(60);
  CQ_PHASE();
  // ---- phase 6: canonical reduction
  if (!cq_reduce<T, L>(g, ra, rb, rc, dbits, !bail)) bail = true;
  if (g_bits<T, L>(g, ra) > 32 * N - 1) bail = true;  // |b| <= a: b fits N limbs signed
  CG_MARK(70);
  g_copy<T, L>(r.a, ra);
  g_copy<T, L>(r.b, rb);
  g_copy<T, L>(r.c, rc);
  return !bail;
}

}  // namespace cofhe
