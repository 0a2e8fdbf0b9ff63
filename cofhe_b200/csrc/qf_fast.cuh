// cofhe_b200 -- register-resident fast path: NUCOMP / NUDUPL with partial reduction and canonical reduction on
// fixed-width limbs (N = limbs of a and b), one form per thread, no local memory, no dynamic indexing.
//
// Same mathematics as qform.cuh (the robust path) -- the 2x4 matrix view of composition with the C and D columns
// recovered through the orthogonality relations -- but organised for the GPU:
//   * Euclid runs as single-precision Lehmer rounds (32-bit leading digits, Jebelean's exactness test), operands
//     kept left-justified by whole limbs so every register index is static;
//   * the exact divisions by a1 are 2-adic (Jebelean) on half-width operands: only the low limbs of the
//     numerators are ever formed;
//   * everything after the partial Euclid is half-width (M limbs) two's-complement/sign-magnitude arithmetic.
// Whatever falls outside the shapes this path is built for (gcd(a1,a2) >= 2^32, a quotient of 31+ bits, tiny `a` /
// huge `c`, a1 divisible by 2^32, ...) is detected and reported through `bail`; the caller then runs the robust
// path on that element.  Both paths end in the unique reduced form, so which one ran is not observable.
#pragma once
#include "fw.cuh"
#include "lehmer.cuh"

// Test hook: a host build may define QF_NOTE to record which guard declined an element (tests/hostsim).
#ifndef QF_NOTE
#define QF_NOTE ((void)0)
#endif
#define QF_BAIL do { QF_NOTE; return false; } while (0)
// Warp reconvergence after a data-dependent loop: the lanes that enter together wait for each other at QF_CONVERGE.
#if defined(__CUDA_ARCH__)
#define QF_CONVERGE_DECL(m) unsigned m = __activemask()
#define QF_CONVERGE(m) __syncwarp(m)
#define QF_ANY(m, p) __any_sync(m, p)
#else
#define QF_ANY(m, p) (p)
#define QF_CONVERGE_DECL(m) ((void)0)
#define QF_CONVERGE(m) ((void)0)
#endif
// Test hook: event counters of a host build (rounds, fallback steps, ...)
#ifndef QF_STAT
#define QF_STAT(k) ((void)0)
#endif

namespace cofhe {

template <int N>
struct FormR {
  u32 a[N];      // > 0
  u32 b[N];      // two's complement
  u32 c[N + 1];  // > 0
};

// floor(a / b) and remainder for a >= b > 0: float reciprocal estimate biased low, then exact fix-up
QF_HD u32 divmod32(u32 a, u32 b, u32 *rem) {
  u32 q = 0, r = a;
  do {
    float fb = (float)b, fr = (float)r, rc;
#if defined(__CUDA_ARCH__)
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(rc) : "f"(fb));
#else
    rc = 1.0f / fb;
#endif
    u32 qe = (u32)(fr * rc * 0.9999995f);
    u64 p = (u64)qe * b;
    if (p > r) { qe = r / b; p = (u64)qe * b; }  // never expected (estimate is biased low)
    if (qe == 0) { qe = 1; p = b; }
    q += qe;
    r -= (u32)p;
  } while (r >= b);
  *rem = r;
  return q;
}

QF_HD float rcp_approx(float x) {
#if defined(__CUDA_ARCH__)
  float r;
  asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
  return r;
#else
  return 1.0f / x;
#endif
}

// A branch-free Euclid step that is switched off never keeps what it computes (the selects at its end), so its divisor
// goes into the reciprocal as it is -- 0 gives inf / NaN, which fail the quotient-size test -- and the quotient is
// converted unguarded.  The device conversion saturates (cvt.rzi.u32.f32: NaN -> 0, too large -> 2^32 - 1); a host build
// spells out the same values, so the simulator runs exactly what the SM runs.
QF_HD u32 hg_quotient(float qf) {
#if defined(__CUDA_ARCH__)
  return (u32)qf;
#else
  return qf >= 4294967296.0f ? 0xffffffffu : (qf >= 0.0f ? (u32)qf : 0u);
#endif
}

// The quotient estimate is q or q - 1: if (rn >= rc) { q++; rn -= rc; } as one 64-bit addition of -rc to (q : rn) -- the
// carry out of the low word is exactly rn >= rc (rc > 0) and lands in q -- and an unsigned minimum that keeps the
// remainder that did not wrap.
QF_HD void hg_fixup(u32 &q, u32 &rn, u32 rc) {
  u64 t = (((u64)q << 32) | rn) + (u64)(u32)(0u - rc);
  u32 lo = (u32)t;
  q = (u32)(t >> 32);
  rn = lo < rn ? lo : rn;
}

// Single-precision Lehmer step: as hgcd2 (lehmer.cuh) on 32-bit leading parts; cofactors stay below 2^30.
//
// This loop is where a composition spends most of its instructions, so it is written for the SM: the quotient comes
// from one reciprocal (float estimate biased low by 2^-20: for q < 2^19 it is q or q-1, one conditional fix-up), the
// float of the divisor is reused (pre-scaled, off the dependency chain) as the next dividend, and two Euclid steps are
// unrolled so that the (previous, current) pairs swap roles instead of moving through registers and the parity of
// Jebelean's test is static.  A quotient of 2^19 or more ends the run; the caller then takes one full-precision step.
//
// EXACT == false (the leading parts are truncations): every accepted step passes Jebelean's test, and the cofactors
// need no overflow check at all -- for the Euclidean sequence of (ah, bh), T_{j+1} r_j + T_j r_{j+1} = ah (same for S
// with bh <= ah), so with the divisor r_j >= 2 the new cofactors are below 2^31 (32-bit arithmetic cannot wrap), and
// an accepted step has T_{j+1} <= r_{j+1} < r_j, hence T_{j+1}^2 < 2^32: accepted cofactors are below 2^16.
// EXACT == true (the operands fit the leading limb: last rounds of a full gcd): no test, cofactors may grow up to
// ah / r_j, so they are formed in 64 bits and the run stops before one reaches 2^30.
#ifndef QF_HG_PAIRS
#define QF_HG_PAIRS 5
#endif
template <bool EXACT>
QF_HD Mat2 hgcd32(u32 ah, u32 bh, u32 min_r) {
  u32 rp = ah, rc = bh;
  u32 Sp = 1, Tp = 0, Sc = 0, Tc = 1;
  const u32 lim = EXACT ? (min_r ? min_r : 1u) : (min_r > 2u ? min_r : 2u);
  const float kLow = 0.99999905f;
  float fprev = (float)rp * kLow;  // dividend estimate, already biased low
  int j = 1;
  bool swapped = false;
  // one Euclid step: (P*) previous, (C*) current; on success the new remainder and cofactors replace the P set
#define QF_HG_STEP(RP, SP, TP, RC, SC, TC, ODD)                                                        \
  {                                                                                                    \
    if (RC < lim) break;                                                                               \
    float fcur = (float)RC;                                                                            \
    float qf = fprev * rcp_approx(fcur);                                                               \
    if (!(qf < 524288.0f)) break;                                                                      \
    u32 q = (u32)qf;                                                                                   \
    u32 rn = RP - q * RC;                                                                              \
    if (rn >= RC) { q++; rn -= RC; }                                                                   \
    u32 Sn, Tn;                                                                                        \
    if (EXACT) {                                                                                       \
      u64 S64 = (u64)q * SC + SP, T64 = (u64)q * TC + TP;                                              \
      if ((S64 | T64) >> 30) break;                                                                    \
      Sn = (u32)S64; Tn = (u32)T64;                                                                    \
    } else {                                                                                           \
      Sn = q * SC + SP; Tn = q * TC + TP;                                                              \
      bool ok = ODD ? (rn >= Tn && RC - rn >= Sn + SC) : (rn >= Sn && RC - rn >= Tn + TC);             \
      if (!ok) break;                                                                                  \
    }                                                                                                  \
    RP = rn; SP = Sn; TP = Tn;                                                                         \
    fprev = fcur * kLow;                                                                               \
    j++;                                                                                               \
  }
  // A truncated run is capped at QF_HG_PAIRS double steps: the lanes of a warp wait for each other at the end of a
  // round, and the number of steps Jebelean's test lets through varies (9 +- 3), so a cap near the mean (measured best at 5 double steps: 3, 4, 6, 8 are 1-9 % slower) keeps
  // nearly all lanes busy for the whole round; any prefix of the Euclidean sequence is a valid round.
  for (int it = 0; EXACT || it < QF_HG_PAIRS; it++) {
    swapped = false;
    QF_HG_STEP(rp, Sp, Tp, rc, Sc, Tc, true)   // index j odd -> even
    swapped = true;
    QF_HG_STEP(rc, Sc, Tc, rp, Sp, Tp, false)  // index j even -> odd
    swapped = false;
  }
#undef QF_HG_STEP
  Mat2 m;
  // after an odd number of accepted steps the sets have swapped roles
  m.S0 = swapped ? Sc : Sp; m.T0 = swapped ? Tc : Tp;
  m.S1 = swapped ? Sp : Sc; m.T1 = swapped ? Tp : Tc;
  m.j = j;
  return m;
}


// hgcd32 without branches (the cooperative path uses the same routine: coop_qf.cuh hgcd32_flat): QF_HG_PAIRS double steps
// are always executed, a step that is not accepted switches the rest off; the accepted steps -- and so the matrix -- are
// those of hgcd32<false>.  Experimental on the register path (option "fast_flat"): no divergence inside the round, at the
// price of predicated work in the lanes that stopped early.
QF_HD Mat2 hgcd32_nb(u32 ah, u32 bh, u32 min_r, bool on) {
  u32 rp = ah, rc = bh;
  u32 Sp = 1, Tp = 0, Sc = 0, Tc = 1;
  const u32 lim = min_r > 2u ? min_r : 2u;
  const float kLow = 0.99999905f;
  float fprev = (float)rp * kLow;
  int j = 1;
  bool go = on;
#define QF_NB_STEP(RP, SP, TP, RC, SC, TC, ODD)                                                        \
  {                                                                                                    \
    bool ok = go && RC >= lim;                                                                         \
    float fcur = (float)RC;                                                                            \
    float qf = fprev * rcp_approx(fcur);                                                               \
    ok = ok && (qf < 524288.0f);                                                                       \
    u32 q = hg_quotient(qf);                                                                           \
    u32 rn = RP - q * RC;                                                                              \
    hg_fixup(q, rn, RC);                                                                               \
    u32 Sn = q * SC + SP, Tn = q * TC + TP;                                                            \
    ok = ODD ? (ok & (rn >= Tn) & (RC - rn >= Sn + SC)) : (ok & (rn >= Sn) & (RC - rn >= Tn + TC));    \
    RP = ok ? rn : RP; SP = ok ? Sn : SP; TP = ok ? Tn : TP;                                           \
    fprev = fcur * kLow; /* after a refused step every later one is off: no select needed */           \
    j += ok ? 1 : 0;                                                                                   \
    go = ok;                                                                                           \
  }
  QF_UNROLL for (int it = 0; it < QF_HG_PAIRS; it++) {
    QF_NB_STEP(rp, Sp, Tp, rc, Sc, Tc, true)
    QF_NB_STEP(rc, Sc, Tc, rp, Sp, Tp, false)
  }
#undef QF_NB_STEP
  const bool swapped = ((j - 1) & 1) != 0;
  Mat2 m;
  m.S0 = swapped ? Sc : Sp; m.T0 = swapped ? Tc : Tp;
  m.S1 = swapped ? Sp : Sc; m.T1 = swapped ? Tp : Tc;
  m.j = j;
  return m;
}

// r = |x*A - y*B| where the sign is known: neg == true means the combination is <= 0.  x, y < 2^30.
// Both products run as carry chains of 32x32+64 multiply-adds (one IMAD.WIDE per limb), then one borrow chain.
template <int N>
QF_HD void fw_lin_sub(u32 *r, u32 x, const u32 *A, u32 y, const u32 *B, bool neg) {
  u32 U[N], V[N];
  u64 cu = 0, cv = 0;
  QF_UNROLL for (int i = 0; i < N; i++) {
    cu = (u64)x * A[i] + (cu >> 32);
    cv = (u64)y * B[i] + (cv >> 32);
    U[i] = (u32)cu;
    V[i] = (u32)cv;
  }
  // the limbs above N cancel (|x*A - y*B| < 2^(32N)); subtract in the order that gives the magnitude
  u32 bo = 0;
  QF_UNROLL for (int i = 0; i < N; i++) {
    u32 p = neg ? V[i] : U[i], q = neg ? U[i] : V[i];
    u64 d = (u64)p - q - bo;
    r[i] = (u32)d;
    bo = (u32)(d >> 63);
  }
}
// r = x*A + y*B, returns the carry-out limb.  x, y < 2^30 (each partial sum fits 64 bits)
template <int N>
QF_HD u32 fw_lin_add(u32 *r, u32 x, const u32 *A, u32 y, const u32 *B) {
  u64 c = 0;
  QF_UNROLL for (int i = 0; i < N; i++) {
    c = (u64)x * A[i] + (u64)y * B[i] + (c >> 32);
    r[i] = (u32)c;
  }
  return (u32)(c >> 32);
}

// Half-extended Euclid on fixed-width magnitudes, same contract as z_hxgcd (lehmer.cuh): in (R0 >= R1, C0, C1),
// runs until the first remainder with bits <= stop_bits, returns the number of Euclidean steps.  `bail` is
// raised when the sequence needs something this path does not do (a 31+ bit quotient, cofactor overflow).
template <int N, int NC, bool FLAT = false>
QF_HD int fw_hxgcd(u32 *R0, u32 *R1, u32 *C0, u32 *C1, int stop_bits, bool &bail) {
  int lsh = fw_lead_zero_limbs<N>(R0);
  if (lsh >= N) { { bail = true; QF_NOTE; } return 0; }
  fw_shl_limbs<N>(R0, lsh);
  fw_shl_limbs<N>(R1, lsh);
  int sb = stop_bits + 32 * lsh;
  int steps = 0;
  // Warp discipline.  The lanes of a warp run different numbers of rounds and, inside a round, different numbers
  // of scalar Euclid steps; left to itself the hardware then executes the multi-limb update of every lane group
  // separately (measured: 5 of 32 lanes active there, half of all issued instructions).  So the lanes that enter
  // together stay together: rounds are counted by a warp vote, lanes that are finished idle inside the loop, and
  // everybody meets again after the scalar loop.
  QF_CONVERGE_DECL(lanes);
  bool done = false;
  // Two passes over the same loop body.  The last rounds of a full gcd run on operands that fit one limb (`exact`: no
  // truncation, no Jebelean test, a long scalar run), and every lane gets there in a different round: taken where they
  // fall, those runs execute a handful of lanes at a time (measured: 6.8 % of all issued instructions at 8 of 32
  // lanes).  So a lane whose operands have come down to one limb waits (`tail` == false) until the others are there or
  // finished, and the exact rounds of the whole warp then run together.  Per lane the sequence of rounds is unchanged.
  bool tail = false;
  // invariant at the top of a round: the top limb of R0 is non-zero (operands are kept left-justified by whole limbs)
  for (int guard = 0;; guard++) {
    int b1 = 0;
    if (!done) {
      // R1 < R0 and R0 fills its top limb: R1 is almost always within the top two limbs
      if (R1[N - 1]) b1 = 32 * N - clz32(R1[N - 1]);
      else if (N >= 2 && R1[N >= 2 ? N - 2 : 0]) b1 = 32 * (N - 1) - clz32(R1[N >= 2 ? N - 2 : 0]);
      else b1 = fw_bits<N>(R1);
      if (b1 <= sb) done = true;  // includes R1 == 0
      else if (guard >= 80 * N) { bail = true; QF_NOTE; done = true; }
    }
    const bool exact = lsh == N - 1;
    const bool act = !done && exact == tail;
    if (!QF_ANY(lanes, act)) {
      if (tail) break;
      tail = true;  // warp-uniform: what is left, if anything, are the one-limb lanes
      continue;
    }
    Mat2 M;
    M.S0 = 1; M.T0 = 0; M.S1 = 0; M.T1 = 1; M.j = 0;
    if (act) {
      int sh = exact ? 0 : clz32(R0[N - 1]);
      u32 ah = R0[N - 1], bh = R1[N - 1];
      if (sh) {
        ah = (ah << sh) | (R0[N >= 2 ? N - 2 : 0] >> (32 - sh));
        bh = (bh << sh) | (R1[N >= 2 ? N - 2 : 0] >> (32 - sh));
      }
      int h = 32 * (N - 1) - sh;
      u32 min_r = sb > h ? (1u << (sb - h)) : 0u;
      if (exact) M = hgcd32<true>(ah, bh, min_r);
      else if (FLAT) M = hgcd32_nb(ah, bh, min_r, true);
      else M = hgcd32<false>(ah, bh, min_r);
      QF_STAT(0);
    }
    QF_CONVERGE(lanes);
    if (act && M.j == 1) {
      QF_STAT(1);
      // one full-precision division step with a quotient of at most 30 bits (rare: about 1 in 100 compositions)
      int b0 = 32 * N - clz32(R0[N - 1]);
      bool ok = b0 - b1 <= 29;
      u32 t[N], cn[NC + 1];
      if (ok) {
        int s2 = clz32(R0[N - 1]);
        u64 A64 = ((u64)R0[N - 1] << 32) | R0[N >= 2 ? N - 2 : 0];
        u64 B64 = ((u64)R1[N - 1] << 32) | R1[N >= 2 ? N - 2 : 0];
        if (s2) {
          u32 a0 = N >= 3 ? R0[N >= 3 ? N - 3 : 0] : 0, b0l = N >= 3 ? R1[N >= 3 ? N - 3 : 0] : 0;
          A64 = (A64 << s2) | (a0 >> (32 - s2));
          B64 = (B64 << s2) | (b0l >> (32 - s2));
        }
        u64 q64 = (B64 == ~(u64)0) ? 0 : A64 / (B64 + 1);
        u32 q = (u32)q64;
        fw_copy<N>(t, R0);
        u32 bo = fw_submul1<N>(t, R1, q);
        ok = bo == 0;  // q underestimates
        int fix = 0;
        while (ok && !fw_lt<N>(t, R1)) {
          fw_sub<N>(t, t, R1);
          q++;
          if (++fix > 8) ok = false;
        }
        if (ok) {
          fw_mul1<NC>(cn, C1, q);
          u32 cy = fw_add<NC>(cn, cn, C0);
          ok = (cn[NC] | cy) == 0;
        }
      }
      if (ok) {
        fw_copy<N>(R0, R1);
        fw_copy<N>(R1, t);
        fw_copy<NC>(C0, C1);
        fw_copy<NC>(C1, cn);
        steps++;
      } else { bail = true; QF_NOTE; done = true; }
    } else if (act) {
      u32 t0[N], t1[N];
      bool odd = M.j & 1;
      // j odd : X_{j-1} = S0*A - T0*B >= 0 , X_j = T1*B - S1*A >= 0 ; j even: signs flipped
      fw_lin_sub<N>(t0, M.S0, R0, M.T0, R1, !odd);
      fw_lin_sub<N>(t1, M.S1, R0, M.T1, R1, odd);
      fw_copy<N>(R0, t0);
      fw_copy<N>(R1, t1);
      u32 c0[NC], c1[NC];
      u32 o0 = fw_lin_add<NC>(c0, M.S0, C0, M.T0, C1);
      u32 o1 = fw_lin_add<NC>(c1, M.S1, C0, M.T1, C1);
      if (o0 | o1) { bail = true; QF_NOTE; done = true; }
      else {
        fw_copy<NC>(C0, c0);
        fw_copy<NC>(C1, c1);
        steps += M.j - 1;
        QF_STAT(3 + (M.j - 1 > 28 ? 28 : M.j - 1));
      }
    }
    // renormalise by whole limbs (a round removes at most ~30 bits, so this runs 0 or 1 times)
    while (act && !done && R0[N - 1] == 0 && lsh < N - 1) {
      QF_UNROLL for (int i = N - 1; i > 0; i--) { R0[i] = R0[i - 1]; R1[i] = R1[i - 1]; }
      R0[0] = 0; R1[0] = 0;
      lsh++; sb += 32;
      QF_STAT(2);
    }
  }
  fw_shr_limbs<N>(R0, lsh);
  fw_shr_limbs<N>(R1, lsh);
  return steps;
}

// exact 32-bit Euclid with one signed cofactor: g = gcd(a, b), y*b == g (mod a); a >= 1
QF_HD u32 xgcd32(u32 a, u32 b, i64 *y) {
  u32 r0 = a, r1 = b;
  i64 y0 = 0, y1 = 1;
  while (r1) {
    u32 r2;
    u32 q = r0 >= r1 ? divmod32(r0, r1, &r2) : (r2 = r0, 0u);
    i64 y2 = y0 - (i64)q * y1;
    r0 = r1; r1 = r2; y0 = y1; y1 = y2;
  }
  *y = y0;
  return r0;
}

// canonical reduction on fixed width: a[N], b two's complement [N+1], c[N+1]; returns false on bail.
template <int N>
QF_HD bool fw_reduce(u32 *a, u32 *b, u32 *c, int dbits) {
  constexpr int W = N + 1;
  for (int it = 0; it < 64; it++) {
    // ---- normalise: -a < b <= a
    int ba = fw_bits<N>(a);
    // c stays within W limbs only while a is not too small
    if (ba == 0 || ba < dbits - 32 * W + 3) QF_BAIL;
    bool bneg = b[W - 1] >> 31;
    u32 bm[W];
    fw_cneg<W>(bm, b, bneg);
    int bb = fw_bits<W>(bm);
    if (bb >= ba) {
      if (bb - ba > 28) QF_BAIL;
      // k = round-ish(b / 2a) from the leading 64 bits, then exact fix-up
      u32 aw[W];
      fw_ext<W, N>(aw, a, false);
      int sh = 32 * W - bb;  // left shift that brings |b|'s msb to the top
      // leading 64 bits of |b| and of a under the same shift (static extraction through a barrel shift)
      u32 tb[W], ta[W];
      fw_copy<W>(tb, bm);
      fw_copy<W>(ta, aw);
      fw_shl_limbs<W>(tb, sh >> 5);
      fw_shl_limbs<W>(ta, sh >> 5);
      u32 tb2[W], ta2[W];
      fw_shl_bits<W>(tb2, tb, sh & 31);
      fw_shl_bits<W>(ta2, ta, sh & 31);
      u64 B64 = ((u64)tb2[W - 1] << 32) | tb2[W - 2];
      u64 A64 = ((u64)ta2[W - 1] << 32) | ta2[W - 2];
      double kd = (double)B64 / (2.0 * (double)A64 + 2.0);
      u32 k = (u32)kd;  // underestimate of floor(|b| / 2a), < 2^29
      // r = |b| - 2a*k  (>= 0)
      u32 a2[W];
      fw_shl_bits<W>(a2, aw, 1);
      u32 r[W];
      fw_copy<W>(r, bm);
      if (k) { if (fw_submul1<W>(r, a2, k)) QF_BAIL; }
      // bring r into (-a, a] for b > 0, [-a, a) for b < 0  (r is treated as signed from here)
      int fix = 0;
      for (;;) {
        bool rneg = r[W - 1] >> 31;
        if (rneg) break;  // can only get here after an overshoot below; not expected
        int cmp = fw_cmp<W>(r, aw);
        bool over = bneg ? cmp >= 0 : cmp > 0;  // r >= a (b<0)  /  r > a (b>0)
        if (!over) break;
        fw_sub<W>(r, r, a2);
        k++;
        if (++fix > 8) QF_BAIL;
      }
      // signed k and new b:  b' = sign * r ; c' = c - ks*(b + b')/2 , with (b + b')/2 = sign*(|b| + r)/2
      u32 hs[W];
      fw_add<W>(hs, bm, r);  // |b| + r  (r may be negative two's complement; sum >= 0)
      fw_shr_bits<W>(hs, hs, 1);
      // c' = c - k*hs  (sign*sign = +)
      if (k) {
        u32 bo = fw_submul1<W>(c, hs, k);
        if (bo) QF_BAIL;
      }
      fw_cneg<W>(b, r, bneg);
    }
    // ---- rho step if a > c
    u32 aw[W];
    fw_ext<W, N>(aw, a, false);
    int cmp = fw_cmp<W>(aw, c);
    if (cmp <= 0) {
      if (cmp == 0 && (b[W - 1] >> 31)) fw_neg<W>(b, b);
      return true;
    }
    if (c[N] != 0) QF_BAIL;  // cannot be: c < a
    u32 t[N];
    fw_copy<N>(t, a);
    fw_copy<N>(a, c);
    fw_ext<W, N>(c, t, false);
    fw_neg<W>(b, b);
  }
  QF_BAIL;
}

// r = f1 o f2 (neg2: f1 o f2^-1; DUPL: f1^2, f2 ignored).  Returns false (bail) when the element must take the
// robust path; r is unspecified then.  dbits = bit length of |Delta|.
//
// The body is a sequence of phases; a phase that declines returns false and the later ones are skipped.  With SYNC the
// warps of a block meet at a block barrier between phases (every thread of the block must then call this function,
// whatever its data): the two Euclid loops are small and cache-resident, but the straight-line multi-limb code between
// them is ~250 KB of SASS that each warp executes once per composition, and the SM's instruction cache serves it to
// all resident warps together only while they stay in step.
#if defined(__CUDA_ARCH__)
#define QF_PHASE() do { if (SYNC) __syncthreads(); } while (0)
#else
#define QF_PHASE() ((void)0)
#endif
template <int N, bool DUPL, bool SYNC = false, bool FLAT = false>
QF_HD bool fq_compose(FormR<N> &r, const FormR<N> &f1, const FormR<N> &f2, bool neg2, int dbits) {
  constexpr int W = N + 1;
  constexpr int M = (N + 1) / 2 + 1;  // half width of the post-Euclid quantities
  // numerators carry extra limbs for the 2-adic shift: a1 = 2^tz * odd with tz < 32*(L-M).  In these non-maximal
  // orders (Delta = -2^(2k+5) N) P(2^j | a) decays only like 2^(-j/2), so tz >= 32 does occur (~2^-17 per step).
  constexpr int L = (M + 2 <= W) ? M + 2 : M + 1;
  bool bail = false, ok = true;

  // ---- operands, larger `a` first
  u32 a1[N], a2[N], b1[W], b2[W], c2[W];
  bool sw = !DUPL && fw_lt<N>(f1.a, f2.a);
  {
    const FormR<N> &p = sw ? f2 : f1, &q = DUPL ? f1 : (sw ? f1 : f2);
    fw_copy<N>(a1, p.a);
    fw_copy<N>(a2, q.a);
    fw_ext<W, N>(b1, p.b, true);
    fw_ext<W, N>(b2, q.b, true);
    fw_copy<W>(c2, q.c);
    if (!DUPL && neg2) { if (sw) fw_neg<W>(b1, b1); else fw_neg<W>(b2, b2); }
  }
  // s = (b1+b2)/2, m = (b2-b1)/2  (two's complement, W limbs)
  u32 s[W], m[W];
  fw_add<W>(s, b1, b2);
  fw_shr_bits<W>(s, s, 1, s[W - 1] >> 31);
  fw_sub<W>(m, b2, s);

  // ---- phase 1: F = gcd(a1, a2) with v*a2 == F (mod a1)   [DUPL: F = a, G = gcd(a, b) = . + y*|b|]
  u32 By[N], Bx[N];
  u32 G = 1;
  u32 R0[N], Cv[N];  // gcd and cofactor of the first Euclid
  int st1 = 0;
  u32 bneg = b1[W - 1] >> 31;
  ok = [&]() -> bool {
    u32 R1[N], C1[N];
    if (!DUPL) {
      fw_copy<N>(R0, a1); fw_copy<N>(R1, a2);
    } else {
      u32 bm[W];
      fw_cneg<W>(bm, b1, bneg);
      if (bm[N] != 0) QF_BAIL;
      if (fw_is_zero<N>(bm)) QF_BAIL;  // b == 0 (ambiguous forms): robust path
      fw_copy<N>(R0, a1); fw_copy<N>(R1, bm);
      // |b| <= a for reduced input; an unreduced input may have |b| > a: robust path
      if (fw_lt<N>(R0, R1)) QF_BAIL;
    }
    fw_zero<N>(Cv); fw_zero<N>(C1); C1[0] = 1;
    st1 = fw_hxgcd<N, N, FLAT>(R0, R1, Cv, C1, 0, bail);
    if (bail) QF_BAIL;
    if (!fw_high_zero<N, 1>(R0)) QF_BAIL;  // gcd >= 2^32: robust path
    return true;
  }();
  QF_PHASE();

  // ---- phase 2: By = a1/G and Bx
  if (ok) ok = [&]() -> bool {
  if (!DUPL) {
    u32 F = R0[0];
    bool vneg = (st1 & 1) == 0;  // sign(C_st) = (-1)^(st+1)
    // t = |v| * |m| mod By, with By = a1/G
    u32 mneg = m[W - 1] >> 31;
    u32 mm[W];
    fw_cneg<W>(mm, m, mneg);
    if (mm[N] != 0) QF_BAIL;
    u32 x0m = 1; bool x0neg = false; u32 y0m = 0; bool y0neg = false;
    u32 x0big[N]; bool x0_is_big = false;
    fw_zero<N>(x0big);
    if (F != 1) {
      // G = gcd(F, s) = x0*F + y0*s
      u32 sneg = s[W - 1] >> 31;
      u32 sm_[W];
      fw_cneg<W>(sm_, s, sneg);
      if (sm_[N] != 0) QF_BAIL;
      u32 srem = fw_mod1<N>(sm_, F);  // |s| mod F
      if (srem == 0) {
        G = F;
      } else {
        i64 y;
        G = xgcd32(F, srem, &y);       // y*|s| == G (mod F)
        if (sneg) y = -y;              // y*s == G (mod F)
        // reduce y into (-F, F) is already true; x0 = (G - y*s)/F  (exact), N+1 limbs signed
        y0neg = y < 0;
        y0m = (u32)(y0neg ? -y : y);
        // |y|*|s| then sign, G - that, / F
        u32 ys[W];
        fw_mul1<N>(ys, sm_, y0m);
        bool ysneg = (y0neg != (sneg != 0)) && !fw_is_zero<W>(ys);
        // num = G - y*s = G - (ysneg ? -ys : ys) = ysneg ? G + ys : G - ys
        u32 num[W], g1[W];
        fw_zero<W>(g1); g1[0] = G;
        bool nneg;
        if (ysneg) { fw_add<W>(num, g1, ys); nneg = false; }
        else {
          if (fw_lt<W>(g1, ys)) { fw_sub<W>(num, ys, g1); nneg = true; }
          else { fw_sub<W>(num, g1, ys); nneg = false; }
        }
        u32 xq[W];
        fw_divexact1<W>(xq, num, F);
        if (xq[N] != 0) QF_BAIL;
        fw_copy<N>(x0big, xq);
        x0neg = nneg;
        x0_is_big = true;
      }
    }
    // By = a1 / G
    if (G == 1) fw_copy<N>(By, a1); else fw_divexact1<N>(By, a1, G);
    // t = (|v|*|m|) mod By
    u32 prod[2 * N], t[N];
    fw_mul<N, N>(prod, Cv, mm);
    if (!fw_mod<2 * N, N>(t, prod, By)) QF_BAIL;
    bool tneg = (vneg != (mneg != 0));
    if (F != 1) {
      // t <- x0 * t (mod By)
      if (x0_is_big) {
        u32 xr[N];
        if (!fw_mod<N, N>(xr, x0big, By)) QF_BAIL;
        fw_mul<N, N>(prod, xr, t);
        if (!fw_mod<2 * N, N>(t, prod, By)) QF_BAIL;
        tneg = tneg != x0neg;
      }
      (void)x0m;
      // u = y0 * (c2 mod By) (mod By)
      if (y0m) {
        u32 cr[N];
        if (!fw_mod<W, N>(cr, c2, By)) QF_BAIL;
        u32 yc[W];
        fw_mul1<N>(yc, cr, y0m);
        u32 u[N];
        if (!fw_mod<W, N>(u, yc, By)) QF_BAIL;
        // Bx = (tneg ? -t : t) + (y0neg ? -u : u)  (mod By)
        // fold signs: bring both to [0, By)
        if (tneg && !fw_is_zero<N>(t)) fw_sub<N>(t, By, t);
        tneg = false;
        if (y0neg && !fw_is_zero<N>(u)) fw_sub<N>(u, By, u);
        u32 cy = fw_add<N>(t, t, u);
        if (cy || !fw_lt<N>(t, By)) fw_sub<N>(t, t, By);
      }
    }
    if (tneg && !fw_is_zero<N>(t)) fw_sub<N>(t, By, t);
    fw_copy<N>(Bx, t);
  } else {
    // NUDUPL: s = b, m = 0, F = a.  G = gcd(a, b) = . + y*b ; Bx = y*c mod (a/G)
    G = R0[0];
    bool yneg = ((st1 & 1) == 0) != (bneg != 0);  // y*|b| == G  ->  y*b == G needs the sign of b
    if (G == 1) fw_copy<N>(By, a1); else fw_divexact1<N>(By, a1, G);
    // Bx = y * c mod By
    u32 cr[N];
    if (!fw_mod<W, N>(cr, c2, By)) QF_BAIL;
    u32 prod[2 * N], t[N];
    fw_mul<N, N>(prod, Cv, cr);
    if (!fw_mod<2 * N, N>(t, prod, By)) QF_BAIL;
    if (yneg && !fw_is_zero<N>(t)) fw_sub<N>(t, By, t);
    fw_copy<N>(Bx, t);
  }
  return true;
  }();
  QF_PHASE();

  // ---- phase 3: partial Euclid on (By, Bx) down to about sqrt(a1/a2) * |D|^(1/4)
  int ba1 = fw_bits<N>(a1), ba2 = fw_bits<N>(a2);
  u32 C0[M], C1[M];
  int st = 0;
  if (ok) {
    int stop = (ba1 - ba2) / 2 + dbits / 4;
    if (stop < 1) stop = 1;
    fw_zero<M>(C0); fw_zero<M>(C1); C1[0] = 1;
    st = fw_hxgcd<N, M, FLAT>(By, Bx, C0, C1, stop, bail);  // (By, Bx) <- (R_{i-1}, R_i)
    if (bail) { QF_NOTE; ok = false; }
  }
  QF_PHASE();

  // ---- phase 4: the C and D columns from the orthogonality relations
  // rows: x' = (Ax, Bxp) with Ax = G*C0 >= 0, Bxp = (-1)^idx R_{i-1};  y' = (Ay, Byp) with Ay = (-1)^(idx+1) G*C1, Byp = R_i
  const int idx = 1 + st;
  const bool ay_neg = (idx & 1) == 0, rx_neg = (idx & 1) != 0;
  u32 Ax[M], Ay[M], Rx[M], Ry[M];
  u32 Cx[M], Cy[M], Dx[M], Dy[M];
  if (ok) ok = [&]() -> bool {
  if (!fw_high_zero<N, M>(By) || !fw_high_zero<N, M>(Bx)) QF_BAIL;
  if ((By[M - 1] | Bx[M - 1]) >> 30) QF_BAIL;
  fw_copy<M>(Rx, By);
  fw_copy<M>(Ry, Bx);
  {
    u32 t[M + 1];
    fw_mul1<M>(t, C0, G); if (t[M] | (t[M - 1] >> 30)) QF_BAIL; fw_copy<M>(Ax, t);
    fw_mul1<M>(t, C1, G); if (t[M] | (t[M - 1] >> 30)) QF_BAIL; fw_copy<M>(Ay, t);
  }
  // size guards for the quotients (they must fit M limbs signed):
  //   |C| <= (a2*|B| + |m|*|A|)/a1 ,  |D| <= (|s|*|B| + c2*|A|)/a1
  {
    int bAx = fw_bits<M>(Ax), bAy = fw_bits<M>(Ay), bRx = fw_bits<M>(Rx), bRy = fw_bits<M>(Ry);
    int bm = fw_sbits<W>(m), bs = fw_sbits<W>(s), bc2 = fw_bits<W>(c2);
    int lim = 32 * M - 3;
    int mxA = bAx > bAy ? bAx : bAy, mxR = bRx > bRy ? bRx : bRy;
    int t1 = ba2 + mxR, t2 = bm + mxA, t3 = bs + mxR, t4 = bc2 + mxA;
    int mx = t1 > t2 ? t1 : t2;
    mx = mx > t3 ? mx : t3;
    mx = mx > t4 ? mx : t4;
    if (mx + 2 - (ba1 - 1) > lim) QF_BAIL;
  }
  // 2-adic exact division by a1 = 2^tz * a1o
  int tz = fw_ctz<N>(a1);
  if (tz > 32 * (L - M) - 1) QF_BAIL;
  int tzl = tz >> 5, tzb = tz & 31;
  u32 a1o[M];
  {
    u32 t[N], t2[N];
    fw_copy<N>(t, a1);
    fw_shr_limbs<N>(t, tzl);
    fw_shr_bits<N>(t2, t, tzb);
    fw_copy<M>(a1o, t2);
  }
  u32 a1inv = inv32(a1o[0]);
  // quotient Q = (p*B - q*A)/a1 for (p, q) = (a2, m) -> C and (s, c2) -> D; B, A given as magnitude + sign
  {
    u32 a2l[L], ml[L], sl[L], c2l[L];
    QF_UNROLL for (int i = 0; i < L; i++) {
      a2l[i] = i < N ? a2[i] : 0;
      ml[i] = m[i < W ? i : W - 1];
      sl[i] = s[i < W ? i : W - 1];
      c2l[i] = i < W ? c2[i] : 0;
    }
    static_assert(L <= W, "half width must not exceed the operand width");
#define QF_ORTHO(QOUT, PL, QL, BMAG, BNEG, AMAG, ANEG)                         \
    {                                                                          \
      u32 n1[L], n2[L];                                                        \
      fw_mul_lo<L, L, M>(n1, PL, BMAG);                                        \
      fw_cneg<L>(n1, n1, BNEG);                                                \
      fw_mul_lo<L, L, M>(n2, QL, AMAG);                                        \
      fw_cneg<L>(n2, n2, ANEG);                                                \
      fw_sub<L>(n1, n1, n2);                                                   \
      u32 sh_[L];                                                              \
      fw_shr_limbs<L>(n1, tzl);                                                \
      fw_shr_bits<L>(sh_, n1, tzb);                                            \
      u32 nm[M];                                                               \
      fw_copy<M>(nm, sh_);                                                     \
      fw_divexact_2adic<M>(QOUT, nm, a1o, a1inv);                              \
    }
    if (!DUPL) QF_ORTHO(Cy, a2l, ml, Ry, false, Ay, ay_neg)
    QF_ORTHO(Dy, sl, c2l, Ry, false, Ay, ay_neg)
    if (!DUPL) QF_ORTHO(Cx, a2l, ml, Rx, rx_neg, Ax, false)
    QF_ORTHO(Dx, sl, c2l, Rx, rx_neg, Ax, false)
#undef QF_ORTHO
  }
  if (DUPL) {
    // C column equals the B column for a square
    fw_cneg<M>(Cy, Ry, false);
    fw_cneg<M>(Cx, Rx, rx_neg);
  }
  return true;
  }();
  QF_PHASE();

  // ---- phase 5: the composed form (two's complement, 2M limbs):
  //   a3 = By*Cy - Ay*Dy ; c3 = Bx*Cx - Ax*Dx ; b3 = (Ax*Dy + Ay*Dx) - (Bx*Cy + By*Cx)
  constexpr int P = 2 * M;
  static_assert(P >= W, "2M >= N+1");
  u32 ra[N], rb[W], rc[W];
  if (ok) ok = [&]() -> bool {
  u32 a3[P], b3[P], c3[P];
  {
    // sign-magnitude of the quotients
    bool ncx = Cx[M - 1] >> 31, ncy = Cy[M - 1] >> 31, ndx = Dx[M - 1] >> 31, ndy = Dy[M - 1] >> 31;
    u32 mcx[M], mcy[M], mdx[M], mdy[M];
    fw_cneg<M>(mcx, Cx, ncx); fw_cneg<M>(mcy, Cy, ncy); fw_cneg<M>(mdx, Dx, ndx); fw_cneg<M>(mdy, Dy, ndy);
    u32 p[P], q[P];
    fw_mul<M, M>(p, Ry, mcy); fw_cneg<P>(p, p, ncy);
    fw_mul<M, M>(q, Ay, mdy); fw_cneg<P>(q, q, ay_neg != ndy);
    fw_sub<P>(a3, p, q);
    fw_mul<M, M>(p, Rx, mcx); fw_cneg<P>(p, p, rx_neg != ncx);
    fw_mul<M, M>(q, Ax, mdx); fw_cneg<P>(q, q, ndx);
    fw_sub<P>(c3, p, q);
    fw_mul<M, M>(p, Ax, mdy); fw_cneg<P>(p, p, ndy);
    fw_mul<M, M>(q, Ay, mdx); fw_cneg<P>(q, q, ay_neg != ndx);
    fw_add<P>(b3, p, q);
    fw_mul<M, M>(p, Rx, mcy); fw_cneg<P>(p, p, rx_neg != ncy);
    fw_sub<P>(b3, b3, p);
    fw_mul<M, M>(p, Ry, mcx); fw_cneg<P>(p, p, ncx);
    fw_sub<P>(b3, b3, p);
  }
  // must fit the working shape: 0 < a3 < 2^(32N-1), 0 < c3 < 2^(32W-1), |b3| < 2^(32W-1)
  if ((a3[P - 1] >> 31) || (c3[P - 1] >> 31)) QF_BAIL;
  if (!fw_high_zero<P, N>(a3) || (a3[N - 1] >> 31) || fw_is_zero<P>(a3)) QF_BAIL;
  if (!fw_high_zero<P, W>(c3) || (c3[W - 1] >> 31)) QF_BAIL;
  {
    u32 fill = (b3[P - 1] >> 31) ? 0xffffffffu : 0u;
    u32 o = 0;
    QF_UNROLL for (int i = W; i < P; i++) o |= b3[i] ^ fill;
    if (o || ((b3[W - 1] >> 31) != (fill & 1))) QF_BAIL;
  }
  fw_copy<N>(ra, a3);
  fw_copy<W>(rb, b3);
  fw_copy<W>(rc, c3);
  return true;
  }();
  QF_PHASE();

  // ---- phase 6: canonical reduction
  if (ok) ok = [&]() -> bool {
  if (!fw_reduce<N>(ra, rb, rc, dbits)) QF_BAIL;
  // b must fit N limbs signed
  {
    u32 fill = (rb[W - 1] >> 31) ? 0xffffffffu : 0u;
    if (rb[N] != fill || ((rb[N - 1] >> 31) != (fill & 1))) QF_BAIL;
  }
  fw_copy<N>(r.a, ra);
  fw_copy<N>(r.b, rb);
  fw_copy<W>(r.c, rc);
  return true;
  }();
  return ok;
}
#undef QF_PHASE

}  // namespace cofhe
