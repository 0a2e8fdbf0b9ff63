// cofhe_b200 -- NUCOMP / NUDUPL with partial reduction and canonical reduction on cooperative integers (coop.cuh): one
// form operation per lane group.  The algorithm is the one of the register path (qf_fast.cuh: van der Poorten's 2x4
// matrix view, single-precision Lehmer rounds with Jebelean's test, C and D columns from the orthogonality relations by
// 2-adic exact divisions, float-estimated canonical reduction) at ONE working width of K = T*L limbs, so the guards and
// their proofs carry over; what changes is where the limbs live.  Results are the unique reduced forms (reference
// thirdparty/bicycl/src/bicycl/qfi.inl:523-548), identical to ClassGroup::nucomp / nucompinv / nudupl (qfi.inl:1441-1471).
// Whatever falls outside the shapes handled here is declined (returns false) and redone by the robust path.
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
// the 64 bits of x below bit position `bb` (x >> (bb - 64), or x << (64 - bb) when bb < 64); bb >= 1
template <int T, int L>
CG_FN u64 g_lead64(const Grp<T> &g, const u32 *x, int bb) {
  int ti = (bb - 1) >> 5, sh = 31 - ((bb - 1) & 31);
  u32 w2 = g_limb<T, L>(g, x, ti), w1 = ti >= 1 ? g_limb<T, L>(g, x, ti >= 1 ? ti - 1 : 0) : 0u,
      w0 = ti >= 2 ? g_limb<T, L>(g, x, ti >= 2 ? ti - 2 : 0) : 0u;
  u64 hi = ((u64)w2 << 32) | w1;
  return sh ? ((hi << sh) | (w0 >> (32 - sh))) : hi;
}

// Half-extended Euclid on cooperative magnitudes, the contract of fw_hxgcd (qf_fast.cuh): in (R0 >= R1, C0, C1), runs
// until the first remainder with bits <= stop_bits, returns the number of Euclidean steps.  The operands are kept
// left-justified by whole limbs, so the leading digits always sit in the top lane's top registers.
template <int T, int L>
CG_FN int g_hxgcd(const Grp<T> &g, u32 *R0, u32 *R1, u32 *C0, u32 *C1, int stop_bits, bool &bail) {
  constexpr int K = T * L;
  static_assert(L >= 3, "the leading digits are read from one lane");
  int bits0 = g_bits<T, L>(g, R0);
  if (bits0 == 0) { bail = true; return 0; }
  int lsh = K - (bits0 + 31) / 32;
  g_shl_limbs<T, L>(g, R0, lsh);
  g_shl_limbs<T, L>(g, R1, lsh);
  int sb = stop_bits + 32 * lsh;
  int steps = 0;
  QF_CONVERGE_DECL(lanes);
  bool done = false;
  for (int guard = 0;; guard++) {
    u32 r0h = 0, r0m = 0, r1h = 0, r1m = 0;
    int b1 = 0;
    if (!done) {
      r0h = g_bcast<T>(g, R0[L - 1], T - 1); r0m = g_bcast<T>(g, R0[L - 2], T - 1);
      r1h = g_bcast<T>(g, R1[L - 1], T - 1); r1m = g_bcast<T>(g, R1[L - 2], T - 1);
      if (r1h) b1 = 32 * K - clz32(r1h);
      else if (r1m) b1 = 32 * (K - 1) - clz32(r1m);
      else b1 = g_bits<T, L>(g, R1);
      if (b1 <= sb) done = true;
      else if (guard >= 80 * K) { bail = true; done = true; }
    }
    if (!QF_ANY(lanes, !done)) break;
    Mat2 M;
    M.S0 = 1; M.T0 = 0; M.S1 = 0; M.T1 = 1; M.j = 0;
    if (!done) {
      bool exact = lsh == K - 1;
      int sh = exact ? 0 : clz32(r0h);
      u32 ah = r0h, bh = r1h;
      if (sh) { ah = (ah << sh) | (r0m >> (32 - sh)); bh = (bh << sh) | (r1m >> (32 - sh)); }
      int h = 32 * (K - 1) - sh;
      u32 min_r = sb > h ? (1u << (sb - h)) : 0u;
      if (exact) M = hgcd32<true>(ah, bh, min_r);
      else M = hgcd32<false>(ah, bh, min_r);
      QF_STAT(0);
    }
    QF_CONVERGE(lanes);
    if (!done && M.j == 1) {
      QF_STAT(1);
      // one full-precision division step with a quotient of at most 30 bits
      int b0 = 32 * K - clz32(r0h);
      bool ok = b0 - b1 <= 29;
      u32 t[L], cn[L];
      if (ok) {
        int s2 = clz32(r0h);
        u64 A64 = ((u64)r0h << 32) | r0m, B64 = ((u64)r1h << 32) | r1m;
        if (s2) {
          u32 a0 = g_bcast<T>(g, R0[L - 3], T - 1), b0l = g_bcast<T>(g, R1[L - 3], T - 1);
          A64 = (A64 << s2) | (a0 >> (32 - s2));
          B64 = (B64 << s2) | (b0l >> (32 - s2));
        }
        u64 q64 = (B64 == ~(u64)0) ? 0 : A64 / (B64 + 1);
        u32 q = (u32)q64;
        g_copy<T, L>(t, R0);
        ok = g_submul1<T, L>(g, t, R1, q) == 0;  // q underestimates
        int fix = 0;
        while (ok && g_cmp<T, L>(g, t, R1) >= 0) {
          g_subm<T, L>(g, t, t, R1);
          q++;
          if (++fix > 8) ok = false;
        }
        if (ok) {
          u32 top = g_mul1<T, L>(g, cn, C1, q);
          u32 cy = g_add<T, L>(g, cn, cn, C0);
          ok = (top | cy) == 0;
        }
      }
      if (ok) {
        g_copy<T, L>(R0, R1); g_copy<T, L>(R1, t);
        g_copy<T, L>(C0, C1); g_copy<T, L>(C1, cn);
        steps++;
      } else { bail = true; done = true; }
    } else if (!done) {
      u32 t0[L], t1[L];
      bool odd = M.j & 1;
      // j odd : X_{j-1} = S0*A - T0*B >= 0 , X_j = T1*B - S1*A >= 0 ; j even: signs flipped
      g_lin_sub<T, L>(g, t0, M.S0, R0, M.T0, R1, !odd);
      g_lin_sub<T, L>(g, t1, M.S1, R0, M.T1, R1, odd);
      g_copy<T, L>(R0, t0); g_copy<T, L>(R1, t1);
      u32 o0 = g_lin_add<T, L>(g, t0, M.S0, C0, M.T0, C1);
      u32 o1 = g_lin_add<T, L>(g, t1, M.S1, C0, M.T1, C1);
      if (o0 | o1) { bail = true; done = true; }
      else {
        g_copy<T, L>(C0, t0); g_copy<T, L>(C1, t1);
        steps += M.j - 1;
      }
    }
    // renormalise by whole limbs (a round removes at most ~30 bits, so this runs 0 or 1 times)
    while (!done && lsh < K - 1 && g_bcast<T>(g, R0[L - 1], T - 1) == 0) {
      g_slide_up<T, L>(g, R0, 0);
      g_slide_up<T, L>(g, R1, 0);
      lsh++; sb += 32;
    }
  }
  g_shr_limbs<T, L>(g, R0, lsh);
  g_shr_limbs<T, L>(g, R1, lsh);
  return steps;
}

// canonical reduction (qfi.inl:523-548): a, c unsigned, b two's complement, all K limbs; false = declined
template <int T, int L>
CG_FN bool cq_reduce(const Grp<T> &g, u32 *a, u32 *b, u32 *c, int dbits) {
  constexpr int K = T * L;
  for (int it = 0; it < 64; it++) {
    // ---- normalise: -a < b <= a
    int ba = g_bits<T, L>(g, a);
    if (ba == 0 || ba > 32 * K - 3 || ba < dbits - 32 * K + 3) return false;
    bool bneg = g_sign<T, L>(g, b);
    u32 bm[L];
    g_cneg<T, L>(g, bm, b, bneg);
    int bb = g_bits<T, L>(g, bm);
    if (bb >= ba) {
      if (bb - ba > 28) return false;
      u64 B64 = g_lead64<T, L>(g, bm, bb), A64 = g_lead64<T, L>(g, a, bb);
      double kd = (double)B64 / (2.0 * (double)A64 + 2.0);
      u32 k = (u32)kd;  // underestimate of floor(|b| / 2a), < 2^29
      u32 a2[L], r[L];
      g_shl_bits<T, L>(g, a2, a, 1);
      g_copy<T, L>(r, bm);
      if (k) { if (g_submul1<T, L>(g, r, a2, k)) return false; }
      // bring r into (-a, a] for b > 0, [-a, a) for b < 0  (r is signed from here)
      int fix = 0;
      for (;;) {
        if (g_sign<T, L>(g, r)) break;
        int cmp = g_cmp<T, L>(g, r, a);
        bool over = bneg ? cmp >= 0 : cmp > 0;
        if (!over) break;
        g_subm<T, L>(g, r, r, a2);
        k++;
        if (++fix > 8) return false;
      }
      // b' = sign * r ; c' = c - k * (|b| + r)/2
      u32 hs[L];
      g_addm<T, L>(g, hs, bm, r);
      g_shr_bits<T, L>(g, hs, hs, 1);
      if (k) { if (g_submul1<T, L>(g, c, hs, k)) return false; }
      g_cneg<T, L>(g, b, r, bneg);
    }
    // ---- rho step if a > c
    int cmp = g_cmp<T, L>(g, a, c);
    if (cmp <= 0) {
      if (cmp == 0 && g_sign<T, L>(g, b)) g_neg<T, L>(g, b, b);
      return true;
    }
    QF_UNROLL for (int i = 0; i < L; i++) { u32 t = a[i]; a[i] = c[i]; c[i] = t; }
    g_neg<T, L>(g, b, b);
  }
  return false;
}

// registers -> scratch slot, product through the streaming multiplier, low K limbs back to registers
template <int T, int L>
CG_FN void cq_mul_lo(const Grp<T> &g, u32 *r, const u32 *a, const u32 *b, int nb) {
  g_sync<T>(g);
  g_to_scr<T, L>(g, scr_a<T, L>(g), a);
  g_to_scr<T, L>(g, scr_b<T, L>(g), b);
  g_mul_s<T, L>(g, scr_a<T, L>(g), scr_b<T, L>(g), nb, scr_out<T, L>(g));
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
// r = (x * y) mod d, x and y unsigned K-limb values with x*y below 2^(32*(K + limbs(y)))
template <int T, int L>
CG_FN bool cq_mulmod(const Grp<T> &g, u32 *r, const u32 *x, const u32 *y, const u32 *d) {
  constexpr int K = T * L;
  int nb = (g_bits<T, L>(g, y) + 31) / 32;
  g_sync<T>(g);
  g_to_scr<T, L>(g, scr_a<T, L>(g), x);
  g_to_scr<T, L>(g, scr_b<T, L>(g), y);
  g_mul_s<T, L>(g, scr_a<T, L>(g), scr_b<T, L>(g), nb, scr_out<T, L>(g));
  int nbr = (nb + L - 1) / L * L;
  g_to_scr<T, L>(g, scr_a<T, L>(g), d);
  bool ok = g_mod_s<T, L>(g, scr_out<T, L>(g), K + nbr, scr_a<T, L>(g), scr_b<T, L>(g));
  g_from_scr<T, L>(g, r, scr_b<T, L>(g));
  return ok;
}

// r = f1 o f2 (neg2: f1 o f2^-1; DUPL: f1^2, f2 ignored).  N = limbs of a reduced a or b.  false = declined.
template <int T, int L, int N, bool DUPL>
CG_FN bool cq_compose(const Grp<T> &g, FormC<L> &r, const FormC<L> &f1, const FormC<L> &f2, bool neg2, int dbits) {
  constexpr int K = T * L;
  constexpr int M = (N + 1) / 2 + 1;  // half width of the post-Euclid quantities
  static_assert(K >= 2 * M && K >= N + 2, "working width too small for this limb count");
  bool bail = false;

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
    if (!DUPL && neg2) { if (sw) g_neg<T, L>(g, b1, b1); else g_neg<T, L>(g, b2, b2); }
  }
  // s = (b1+b2)/2, m = (b2-b1)/2  (two's complement)
  u32 s[L], m[L];
  g_addm<T, L>(g, s, b1, b2);
  g_shr_bits<T, L>(g, s, s, 1, g_sign<T, L>(g, s) ? 0xffffffffu : 0u);
  g_subm<T, L>(g, m, b2, s);

  // ---- phase 1: F = gcd(a1, a2) with v*a2 == F (mod a1)   [DUPL: F = a, G = gcd(a, b) = . + y*|b|]
  u32 By[L], Bx[L], Cv[L];
  u32 G = 1;
  int st1 = 0;
  const bool bneg = g_sign<T, L>(g, b1);
  {
    u32 R1[L], C1[L];
    g_copy<T, L>(By, a1);  // By doubles as R0 of the first Euclid
    if (!DUPL) g_copy<T, L>(R1, a2);
    else {
      g_cneg<T, L>(g, R1, b1, bneg);
      if (g_is_zero<T, L>(g, R1)) return false;     // b == 0 (ambiguous forms): robust path
      if (g_cmp<T, L>(g, By, R1) < 0) return false;  // unreduced input
    }
    g_zero<T, L>(Cv);
    g_set_u32<T, L>(g, C1, 1);
    st1 = g_hxgcd<T, L>(g, By, R1, Cv, C1, 0, bail);
    if (bail) return false;
    if (g_bits<T, L>(g, By) > 32) return false;  // gcd >= 2^32: robust path
  }
  const u32 F = g_bcast<T>(g, By[0], 0);

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
    if (F != 1) {
      // G = gcd(F, s) = x0*F + y0*s
      bool sneg = g_sign<T, L>(g, s);
      u32 sm_[L];
      g_cneg<T, L>(g, sm_, s, sneg);
      u32 srem = g_mod1<T, L>(g, sm_, F);
      if (srem == 0) G = F;
      else {
        i64 y;
        G = xgcd32(F, srem, &y);  // y*|s| == G (mod F)
        if (sneg) y = -y;
        y0neg = y < 0;
        y0m = (u32)(y0neg ? -y : y);
        // x0 = (G - y*s)/F  (exact)
        u32 ys[L], g1[L], num[L];
        if (g_mul1<T, L>(g, ys, sm_, y0m)) return false;
        bool ysneg = (y0neg != sneg) && !g_is_zero<T, L>(g, ys);
        g_set_u32<T, L>(g, g1, G);
        bool nneg = false;
        if (ysneg) g_addm<T, L>(g, num, g1, ys);
        else if (g_cmp<T, L>(g, g1, ys) < 0) { g_subm<T, L>(g, num, ys, g1); nneg = true; }
        else g_subm<T, L>(g, num, g1, ys);
        g_divexact1<T, L>(g, x0big, num, F);
        x0neg = nneg;
        x0_is_big = true;
      }
    }
    if (G == 1) g_copy<T, L>(By, a1); else g_divexact1<T, L>(g, By, a1, G);
    // t = (|v|*|m|) mod By
    u32 t[L];
    if (!cq_mulmod<T, L>(g, t, Cv, mm, By)) return false;
    bool tneg = vneg != mneg;
    if (F != 1) {
      if (x0_is_big) {
        // t <- x0 * t (mod By)
        u32 xr[L];
        if (!cq_mod<T, L>(g, xr, x0big, By)) return false;
        if (!cq_mulmod<T, L>(g, t, xr, t, By)) return false;
        tneg = tneg != x0neg;
      }
      if (y0m) {
        // u = y0 * (c2 mod By) (mod By)
        u32 cr[L], u[L];
        if (!cq_mod<T, L>(g, cr, c2, By)) return false;
        if (g_mul1<T, L>(g, cr, cr, y0m)) return false;
        if (!cq_mod<T, L>(g, u, cr, By)) return false;
        // Bx = (tneg ? -t : t) + (y0neg ? -u : u)  (mod By): bring both to [0, By)
        if (tneg && !g_is_zero<T, L>(g, t)) g_subm<T, L>(g, t, By, t);
        tneg = false;
        if (y0neg && !g_is_zero<T, L>(g, u)) g_subm<T, L>(g, u, By, u);
        u32 cy = g_add<T, L>(g, t, t, u);
        if (cy || g_cmp<T, L>(g, t, By) >= 0) g_subm<T, L>(g, t, t, By);
      }
    }
    if (tneg && !g_is_zero<T, L>(g, t)) g_subm<T, L>(g, t, By, t);
    g_copy<T, L>(Bx, t);
  } else {
    // NUDUPL: s = b, m = 0, F = a.  G = gcd(a, b) = . + y*b ; Bx = y*c mod (a/G)
    G = F;
    bool yneg = ((st1 & 1) == 0) != bneg;
    if (G == 1) g_copy<T, L>(By, a1); else g_divexact1<T, L>(g, By, a1, G);
    u32 cr[L], t[L];
    if (!cq_mod<T, L>(g, cr, c2, By)) return false;
    if (!cq_mulmod<T, L>(g, t, Cv, cr, By)) return false;
    if (yneg && !g_is_zero<T, L>(g, t)) g_subm<T, L>(g, t, By, t);
    g_copy<T, L>(Bx, t);
  }

  // ---- phase 3: partial Euclid on (By, Bx) down to about sqrt(a1/a2) * |D|^(1/4)
  const int ba1 = g_bits<T, L>(g, a1), ba2 = g_bits<T, L>(g, a2);
  u32 C0[L], C1[L];
  int st = 0;
  {
    int stop = (ba1 - ba2) / 2 + dbits / 4;
    if (stop < 1) stop = 1;
    g_zero<T, L>(C0);
    g_set_u32<T, L>(g, C1, 1);
    st = g_hxgcd<T, L>(g, By, Bx, C0, C1, stop, bail);  // (By, Bx) <- (R_{i-1}, R_i)
    if (bail) return false;
  }

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
    if (bRx > lim || bRy > lim) return false;
    if (g_mul1<T, L>(g, Ax, C0, G)) return false;
    if (g_mul1<T, L>(g, Ay, C1, G)) return false;
    int bAx = g_bits<T, L>(g, Ax), bAy = g_bits<T, L>(g, Ay);
    if (bAx > lim || bAy > lim) return false;
    // size guards for the quotients (they must fit M limbs signed):
    //   |C| <= (a2*|B| + |m|*|A|)/a1 ,  |D| <= (|s|*|B| + c2*|A|)/a1
    int bm = g_sbits<T, L>(g, m), bs = g_sbits<T, L>(g, s), bc2 = g_bits<T, L>(g, c2);
    int mxA = bAx > bAy ? bAx : bAy, mxR = bRx > bRy ? bRx : bRy;
    int t1 = ba2 + mxR, t2 = bm + mxA, t3 = bs + mxR, t4 = bc2 + mxA;
    int mx = t1 > t2 ? t1 : t2;
    mx = mx > t3 ? mx : t3;
    mx = mx > t4 ? mx : t4;
    if (mx + 2 - (ba1 - 1) > 32 * M - 3) return false;
    // 2-adic exact division by a1 = 2^tz * a1o: only the low limbs of the numerators are formed
    int tz = g_ctz<T, L>(g, a1);
    if (tz > 32 * (K - M) - 1) return false;
    u32 a1o[L];
    g_copy<T, L>(a1o, a1);
    g_shr_limbs<T, L>(g, a1o, tz >> 5);
    g_shr_bits<T, L>(g, a1o, a1o, tz & 31);
    // quotient Q = (p*B - q*A)/a1 for (p, q) = (a2, m) -> C and (s, c2) -> D; B, A given as magnitude + sign
    auto ortho = [&](u32 *Q, const u32 *P, const u32 *Qn, const u32 *Bmag, bool Bneg, const u32 *Amag, bool Aneg) {
      u32 n1[L], n2[L];
      cq_mul_lo<T, L>(g, n1, P, Bmag, M);
      if (Bneg) g_neg<T, L>(g, n1, n1);
      cq_mul_lo<T, L>(g, n2, Qn, Amag, M);
      if (Aneg) g_neg<T, L>(g, n2, n2);
      g_subm<T, L>(g, n1, n1, n2);
      g_shr_limbs<T, L>(g, n1, tz >> 5);
      g_shr_bits<T, L>(g, n1, n1, tz & 31);
      g_sync<T>(g);
      g_to_scr<T, L>(g, scr_a<T, L>(g), n1);
      g_to_scr<T, L>(g, scr_b<T, L>(g), a1o);
      g_divexact2_s<T, L>(g, scr_a<T, L>(g), scr_b<T, L>(g), M, scr_out<T, L>(g));
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

  // ---- phase 5: the composed form (two's complement):
  //   a3 = By*Cy - Ay*Dy ; c3 = Bx*Cx - Ax*Dx ; b3 = (Ax*Dy + Ay*Dx) - (Bx*Cy + By*Cx)
  u32 ra[L], rb[L], rc[L];
  {
    bool ncx = g_sign<T, L>(g, Cx), ncy = g_sign<T, L>(g, Cy), ndx = g_sign<T, L>(g, Dx), ndy = g_sign<T, L>(g, Dy);
    g_cneg<T, L>(g, Cx, Cx, ncx); g_cneg<T, L>(g, Cy, Cy, ncy); g_cneg<T, L>(g, Dx, Dx, ndx); g_cneg<T, L>(g, Dy, Dy, ndy);
    u32 p[L];
    auto term = [&](u32 *acc, const u32 *x, const u32 *y, bool neg, bool first) {
      cq_mul_lo<T, L>(g, p, x, y, M);
      if (first) g_cneg<T, L>(g, acc, p, neg);
      else if (neg) g_subm<T, L>(g, acc, acc, p);
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
  if (g_sign<T, L>(g, ra) || g_sign<T, L>(g, rc)) return false;
  {
    int b3a = g_bits<T, L>(g, ra), b3c = g_bits<T, L>(g, rc);
    if (b3a == 0 || b3a > 32 * N - 1 || b3c == 0 || b3c > 32 * K - 3) return false;
  }

  // ---- phase 6: canonical reduction
  if (!cq_reduce<T, L>(g, ra, rb, rc, dbits)) return false;
  if (g_bits<T, L>(g, ra) > 32 * N - 1) return false;  // |b| <= a: b fits N limbs signed
  g_copy<T, L>(r.a, ra);
  g_copy<T, L>(r.b, rb);
  g_copy<T, L>(r.c, rc);
  return true;
}

}  // namespace cofhe
