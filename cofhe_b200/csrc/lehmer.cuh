// cofhe_b200 -- Lehmer (leading-digit) Euclid: the scalar 2x2 step shared by the generic and the fixed-width
// paths, and the generic half-extended / partial GCD on dynamic-length integers.
//
// Replaces, with a different algorithmic organisation, what the reference gets from mpz_gcdext
// (thirdparty/bicycl/src/bicycl/gmp_extras.inl:1044-1048) and Mpz::partial_euclid (gmp_extras.inl:1231-1392,
// which drives GMP's internal mpn_hgcd2).  Cofactor conventions do not leak into results: every form
// operation ends in the canonical reduction.
#pragma once
#include "bn.cuh"

namespace cofhe {

// Cofactor magnitudes of two consecutive remainders X_{j-1}, X_j of the Euclidean sequence started at
// (X_0, X_1) = (A, B):  X_i = s_i*A + t_i*B with sign(s_i) = (-1)^i, sign(t_i) = (-1)^(i+1).
// j == 1 means "no step was possible" (identity).
struct Mat2 {
  u32 S0, T0, S1, T1;
  int j;
};

QF_HD u64 udiv64(u64 a, u64 b) {
  // small quotients dominate Euclid's sequence (q = 1: 41.5 %, 2: 17 %, 3: 9.3 % ...)
  u64 d = a - b;
  if (d < b) return 1;
  d -= b;
  if (d < b) return 2;
  d -= b;
  if (d < b) return 3;
  return a / b;
}

// Euclid on the 64-bit leading parts (ah >= bh).  With exact == false the inputs are truncations
// (A = ah*2^h + lo, 0 <= lo < 2^h, same for B) and a step is kept only when
//     j even:  r_j >= T_j  and  r_{j-1} - r_j >= S_j + S_{j-1}
//     j odd :  r_j >= S_j  and  r_{j-1} - r_j >= T_j + T_{j-1}
// which guarantees 0 <= X_j < X_{j-1} for every completion of the low parts (Jebelean's condition).
// Stepping also stops before a step whose divisor is already below min_r (partial GCD bound) and before
// any cofactor would leave 32 bits.
QF_HD Mat2 hgcd2(u64 ah, u64 bh, bool exact, u64 min_r) {
  u64 r0 = ah, r1 = bh;
  u64 S0 = 1, T0 = 0, S1 = 0, T1 = 1;
  int j = 1;
  for (;;) {
    if (r1 == 0 || r1 < min_r) break;
    u64 q = udiv64(r0, r1);
    if (q >> 32) break;
    u64 r2 = r0 - q * r1;
    u64 S2 = S0 + q * S1, T2 = T0 + q * T1;
    if ((S2 | T2) >> 32) break;
    if (!exact) {
      bool ok;
      if (j & 1) ok = r2 >= T2 && (r1 - r2) >= S2 + S1;  // new index j+1 is even
      else ok = r2 >= S2 && (r1 - r2) >= T2 + T1;        // new index j+1 is odd
      if (!ok) break;
    }
    r0 = r1; r1 = r2; S0 = S1; S1 = S2; T0 = T1; T1 = T2;
    j++;
  }
  Mat2 m;
  m.S0 = (u32)S0; m.T0 = (u32)T0; m.S1 = (u32)S1; m.T1 = (u32)T1; m.j = j;
  return m;
}

// r = x*A - y*B, known to be >= 0
template <int CAP>
QF_HDN void z_lin_sub(Err &e, Z<CAP> &r, u32 x, const Z<CAP> &A, u32 y, const Z<CAP> &B) {
  int n = A.n > B.n ? A.n : B.n;
  u64 cp = 0, cq = 0;
  i64 br = 0;
  u32 out[CAP + 1];
  for (int i = 0; i <= n && i <= CAP; i++) {
    u64 p = (u64)x * (i < A.n ? A.d[i] : 0) + cp;
    cp = p >> 32;
    u64 q = (u64)y * (i < B.n ? B.d[i] : 0) + cq;
    cq = q >> 32;
    i64 t = (i64)(u32)p - (i64)(u32)q + br;
    out[i] = (u32)t;
    br = t >> 32;
  }
  if (br != 0) e.code = 3;  // negative: violates the caller's guarantee
  int m = bn_norm(out, (n + 1 <= CAP + 1) ? n + 1 : CAP + 1);
  if (m > CAP) { e.code = 1; m = 0; }
  for (int i = 0; i < m; i++) r.d[i] = out[i];
  r.n = m; r.neg = 0;
}
// r = x*A + y*B (magnitudes)
template <int CAP>
QF_HDN void z_lin_add(Err &e, Z<CAP> &r, u32 x, const Z<CAP> &A, u32 y, const Z<CAP> &B) {
  int n = A.n > B.n ? A.n : B.n;
  u64 cp = 0, cq = 0, cs = 0;
  u32 out[CAP + 2];
  for (int i = 0; i <= n + 1 && i <= CAP + 1; i++) {
    u64 p = (u64)x * (i < A.n ? A.d[i] : 0) + cp;
    cp = p >> 32;
    u64 q = (u64)y * (i < B.n ? B.d[i] : 0) + cq;
    cq = q >> 32;
    u64 sum = (p & 0xffffffffu) + (q & 0xffffffffu) + cs;
    out[i] = (u32)sum;
    cs = sum >> 32;
  }
  int m = bn_norm(out, (n + 2 <= CAP + 2) ? n + 2 : CAP + 2);
  if (m > CAP) { e.code = 1; m = 0; }
  for (int i = 0; i < m; i++) r.d[i] = out[i];
  r.n = m; r.neg = 0;
}

// Half-extended Euclid on magnitudes.  In: R0 >= R1 >= 0 and cofactor magnitudes (C0, C1) (callers start
// from (0, 1)).  Runs the Euclidean sequence R_{i+1} = R_{i-1} - q_i R_i, |C_{i+1}| = |C_{i-1}| + q_i |C_i|
// and stops at the first i with bits(R_i) <= stop_bits (stop_bits == 0: until R_i == 0).  On return
// (R0, R1) = (R_{i-1}, R_i), (C0, C1) = (|C_{i-1}|, |C_i|); the return value is the number of steps taken, so
// with the initial pair at indices (0, 1) the final index is i = 1 + steps and sign(C_i) = (-1)^(i+1).
template <int CAP>
QF_HDN int z_hxgcd(Err &e, Z<CAP> &R0, Z<CAP> &R1, Z<CAP> &C0, Z<CAP> &C1, int stop_bits) {
  int steps = 0;
  Z<CAP> t0, t1;
  for (;;) {
    if (e.code) break;
    if (R1.n == 0 || z_bits(R1) <= stop_bits) break;
    int n = R0.n;
    u64 ah, bh, min_r;
    bool exact;
    if (n <= 2) {
      exact = true;
      ah = R0.d[0] | (n > 1 ? (u64)R0.d[1] << 32 : 0);
      bh = (R1.n > 0 ? R1.d[0] : 0) | (R1.n > 1 ? (u64)R1.d[1] << 32 : 0);
      min_r = 1ull << stop_bits;  // stop_bits < bits(R1) <= 64
    } else {
      exact = false;
      int sh = clz32(R0.d[n - 1]);
      u32 a2 = R0.d[n - 1], a1 = R0.d[n - 2], a0 = R0.d[n - 3];
      u32 b2 = R1.n > n - 1 ? R1.d[n - 1] : 0, b1 = R1.n > n - 2 ? R1.d[n - 2] : 0, b0 = R1.n > n - 3 ? R1.d[n - 3] : 0;
      ah = ((u64)a2 << 32) | a1;
      bh = ((u64)b2 << 32) | b1;
      if (sh) { ah = (ah << sh) | (a0 >> (32 - sh)); bh = (bh << sh) | (b0 >> (32 - sh)); }
      int h = 32 * (n - 2) - sh;
      min_r = stop_bits > h ? (1ull << (stop_bits - h)) : 0;
    }
    Mat2 M = hgcd2(ah, bh, exact, min_r);
    if (M.j == 1) {
      // one full-precision division step
      Z<CAP> q;
      z_tdiv_qr<CAP>(e, q, t0, R0, R1);
      z_copy(e, R0, R1);
      z_copy(e, R1, t0);
      z_mul(e, t1, q, C1);
      z_add(e, t1, t1, C0);
      z_copy(e, C0, C1);
      z_copy(e, C1, t1);
      steps++;
    } else {
      if (M.j & 1) {
        z_lin_sub(e, t0, M.S0, R0, M.T0, R1);
        z_lin_sub(e, t1, M.T1, R1, M.S1, R0);
      } else {
        z_lin_sub(e, t0, M.T0, R1, M.S0, R0);
        z_lin_sub(e, t1, M.S1, R0, M.T1, R1);
      }
      z_copy(e, R0, t0);
      z_copy(e, R1, t1);
      z_lin_add(e, t0, M.S0, C0, M.T0, C1);
      z_lin_add(e, t1, M.S1, C0, M.T1, C1);
      z_copy(e, C0, t0);
      z_copy(e, C1, t1);
      steps += M.j - 1;
    }
  }
  return steps;
}

}  // namespace cofhe
