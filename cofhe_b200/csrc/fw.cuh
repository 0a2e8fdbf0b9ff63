// cofhe_b200 -- fixed-width multi-precision primitives for the register-resident fast path.
//
// Every array length is a template constant and every loop is fully unrolled with compile-time indices, so
// the limbs live in registers (no local-memory traffic, no dynamic indexing).  Arithmetic is written on 64-bit
// intermediates; nvcc lowers `(u64)a * b + c` to IMAD.WIDE.U32 and the carry handling to IADD3/IADD3.X chains.
// Values are little-endian u32 limbs; signed values are two's complement over their width.
#pragma once
#include "bn.cuh"

namespace cofhe {

#if defined(__CUDACC__)
#define QF_UNROLL _Pragma("unroll")
#define QF_NOUNROLL _Pragma("unroll 1")
#else
#define QF_UNROLL
#define QF_NOUNROLL
#endif

template <int N>
QF_HD void fw_zero(u32 *r) {
  QF_UNROLL for (int i = 0; i < N; i++) r[i] = 0;
}
template <int N>
QF_HD void fw_copy(u32 *r, const u32 *a) {
  QF_UNROLL for (int i = 0; i < N; i++) r[i] = a[i];
}
// r[NR] = a[NA] zero- or sign-extended / truncated
template <int NR, int NA>
QF_HD void fw_ext(u32 *r, const u32 *a, bool sign) {
  u32 fill = (sign && (a[NA - 1] >> 31)) ? 0xffffffffu : 0u;
  QF_UNROLL for (int i = 0; i < NR; i++) r[i] = i < NA ? a[i] : fill;
}
template <int N>
QF_HD bool fw_is_zero(const u32 *a) {
  u32 o = 0;
  QF_UNROLL for (int i = 0; i < N; i++) o |= a[i];
  return o == 0;
}
// limbs [LO, N) all zero?
template <int N, int LO>
QF_HD bool fw_high_zero(const u32 *a) {
  u32 o = 0;
  QF_UNROLL for (int i = LO; i < N; i++) o |= a[i];
  return o == 0;
}
template <int N>
QF_HD u32 fw_add(u32 *r, const u32 *a, const u32 *b) {
  u64 c = 0;
  QF_UNROLL for (int i = 0; i < N; i++) { c += (u64)a[i] + b[i]; r[i] = (u32)c; c >>= 32; }
  return (u32)c;
}
// r = a - b, returns borrow (1 if a < b)
template <int N>
QF_HD u32 fw_sub(u32 *r, const u32 *a, const u32 *b) {
  i64 c = 0;
  QF_UNROLL for (int i = 0; i < N; i++) { c += (i64)a[i] - (i64)b[i]; r[i] = (u32)c; c >>= 32; }
  return (u32)(c & 1);
}
template <int N>
QF_HD void fw_neg(u32 *r, const u32 *a) {
  u64 c = 1;
  QF_UNROLL for (int i = 0; i < N; i++) { c += (u32)~a[i]; r[i] = (u32)c; c >>= 32; }
}
// conditional negate (two's complement) without branching
template <int N>
QF_HD void fw_cneg(u32 *r, const u32 *a, bool neg) {
  u32 m = neg ? 0xffffffffu : 0u;
  u64 c = neg ? 1 : 0;
  QF_UNROLL for (int i = 0; i < N; i++) { c += (u32)(a[i] ^ m); r[i] = (u32)c; c >>= 32; }
}
// unsigned compare: -1, 0, 1
template <int N>
QF_HD int fw_cmp(const u32 *a, const u32 *b) {
  int r = 0;
  QF_UNROLL for (int i = 0; i < N; i++) r = a[i] != b[i] ? (a[i] < b[i] ? -1 : 1) : r;
  return r;
}
template <int N>
QF_HD bool fw_lt(const u32 *a, const u32 *b) {  // a < b via borrow chain
  i64 c = 0;
  QF_UNROLL for (int i = 0; i < N; i++) { c += (i64)a[i] - (i64)b[i]; c >>= 32; }
  return c != 0;
}
// bit length of an unsigned value
template <int N>
QF_HD int fw_bits(const u32 *a) {
  int r = 0;
  QF_UNROLL for (int i = 0; i < N; i++) r = a[i] ? 32 * i + 32 - clz32(a[i]) : r;
  return r;
}
// bit length of |a| for two's complement a (approximation safe for bounds: bits of a if a>=0, of ~a if a<0, +0)
template <int N>
QF_HD int fw_sbits(const u32 *a) {
  u32 m = (a[N - 1] >> 31) ? 0xffffffffu : 0u;
  int r = 0;
  QF_UNROLL for (int i = 0; i < N; i++) { u32 w = a[i] ^ m; r = w ? 32 * i + 32 - clz32(w) : r; }
  return r + (m ? 1 : 0);
}
template <int N>
QF_HD int fw_ctz(const u32 *a) {  // a != 0
  int r = 32 * N;
  QF_UNROLL for (int i = N - 1; i >= 0; i--) r = a[i] ? 32 * i + ctz32(a[i]) : r;
  return r;
}
// full product r[NA+NB] = a[NA] * b[NB] (unsigned)
template <int NA, int NB>
QF_HD void fw_mul(u32 *r, const u32 *a, const u32 *b) {
  QF_UNROLL for (int i = 0; i < NA + NB; i++) r[i] = 0;
  QF_UNROLL for (int i = 0; i < NA; i++) {
    u64 c = 0;
    QF_UNROLL for (int j = 0; j < NB; j++) { c += (u64)a[i] * b[j] + r[i + j]; r[i + j] = (u32)c; c >>= 32; }
    r[i + NB] = (u32)c;
  }
}
// low NR limbs of a[NA] * b[NB] (wraps mod 2^(32 NR): exact for two's complement operands extended to NR limbs)
template <int NR, int NA, int NB>
QF_HD void fw_mul_lo(u32 *r, const u32 *a, const u32 *b) {
  QF_UNROLL for (int i = 0; i < NR; i++) r[i] = 0;
  QF_UNROLL for (int i = 0; i < NA; i++) {
    if (i < NR) {
      u64 c = 0;
      QF_UNROLL for (int j = 0; j < NB; j++) {
        if (i + j < NR) { c += (u64)a[i] * b[j] + r[i + j]; r[i + j] = (u32)c; c >>= 32; }
      }
      if (i + NB < NR) r[i + NB] = (u32)c;
    }
  }
}
// r[N+1] = a[N] * w
template <int N>
QF_HD void fw_mul1(u32 *r, const u32 *a, u32 w) {
  u64 c = 0;
  QF_UNROLL for (int i = 0; i < N; i++) { c += (u64)a[i] * w; r[i] = (u32)c; c >>= 32; }
  r[N] = (u32)c;
}
// r[N] -= a[N] * w, returns the borrow-out limb (what would have to be subtracted from r[N])
template <int N>
QF_HD u32 fw_submul1(u32 *r, const u32 *a, u32 w) {
  u64 c = 0;
  QF_UNROLL for (int i = 0; i < N; i++) {
    u64 p = (u64)a[i] * w + c;
    u32 lo = (u32)p;
    c = (p >> 32) + (r[i] < lo ? 1 : 0);
    r[i] -= lo;
  }
  return (u32)c;
}
// shifts by 0 <= s < 32 bits
template <int N>
QF_HD void fw_shl_bits(u32 *r, const u32 *a, int s) {
  if (s == 0) { fw_copy<N>(r, a); return; }
  QF_UNROLL for (int i = N - 1; i > 0; i--) r[i] = (a[i] << s) | (a[i - 1] >> (32 - s));
  r[0] = a[0] << s;
}
template <int N>
QF_HD void fw_shr_bits(u32 *r, const u32 *a, int s, u32 fill = 0) {
  if (s == 0) { fw_copy<N>(r, a); return; }
  QF_UNROLL for (int i = 0; i < N - 1; i++) r[i] = (a[i] >> s) | (a[i + 1] << (32 - s));
  r[N - 1] = (a[N - 1] >> s) | (fill << (32 - s));
}
// shift left/right by k whole limbs, 0 <= k < N, k dynamic: log-depth barrel of selects (static indices only)
template <int N>
QF_HD void fw_shl_limbs(u32 *a, int k) {
  QF_UNROLL for (int st = 1; st < N; st <<= 1) {
    if (k & st) {
      QF_UNROLL for (int i = N - 1; i >= 0; i--) a[i] = i >= st ? a[i - st] : 0;
    }
  }
}
template <int N>
QF_HD void fw_shr_limbs(u32 *a, int k) {
  QF_UNROLL for (int st = 1; st < N; st <<= 1) {
    if (k & st) {
      QF_UNROLL for (int i = 0; i < N; i++) a[i] = i + st < N ? a[i + st] : 0;
    }
  }
}
// number of leading zero limbs
template <int N>
QF_HD int fw_lead_zero_limbs(const u32 *a) {
  int top = -1;
  QF_UNROLL for (int i = 0; i < N; i++) top = a[i] ? i : top;
  return N - 1 - top;
}

// inverse of odd x modulo 2^32 (Newton, 5 doublings from 3 correct bits)
QF_HD u32 inv32(u32 x) {
  u32 y = x * 3u ^ 2u;  // 5 bits
  y *= 2u - x * y;      // 10
  y *= 2u - x * y;      // 20
  y *= 2u - x * y;      // 40
  return y;
}
// q[M] = num[M] / d[M] in Z/2^(32M) (d odd): Jebelean's exact division, least significant limb first.
// num is destroyed.
template <int M>
QF_HD void fw_divexact_2adic(u32 *q, u32 *num, const u32 *d, u32 dinv) {
  QF_UNROLL for (int i = 0; i < M; i++) {
    u32 qi = num[i] * dinv;
    q[i] = qi;
    // num[i..M) -= qi * d[0..M-i)
    u64 c = 0;
    QF_UNROLL for (int j = 0; j < M; j++) {
      if (i + j < M) {
        u64 p = (u64)qi * d[j] + c;
        u32 lo = (u32)p;
        c = (p >> 32) + (num[i + j] < lo ? 1 : 0);
        num[i + j] -= lo;
      }
    }
  }
}

// (qhat, rem) of the 64-by-32 division (u1:u0) / d for normalised d (top bit set) and u1 < d, with a precomputed
// reciprocal v = floor((2^64 - 1) / d) - 2^32  (Moller & Granlund, "Improved division by invariant integers")
QF_HD u32 recip32(u32 d) {
#if defined(__CUDA_ARCH__)
  // floor((2^64 - 1) / d) - 2^32 = floor(t / d) with t = 2^64 - 1 - d * 2^32 < d * 2^32 (d is normalised).  One
  // double-precision division (relative error 2^-52 on a quotient below 2^32: off by at most one) and an exact
  // fix-up, instead of the 64-bit integer division subroutine.
  u64 t = ((u64)(~d) << 32) | 0xffffffffu;
  double q = (double)t / (double)d;
  u32 v = q >= 4294967295.0 ? 0xffffffffu : (u32)q;
  u64 p = (u64)v * d;
  if (p > t) v--;
  else if (t - p >= d) v++;
  return v;
#else
  return (u32)(~(u64)0 / d - ((u64)1 << 32));
#endif
}
QF_HD u32 div_2by1(u32 u1, u32 u0, u32 d, u32 v, u32 *rem) {
  u64 q = (u64)v * u1 + (((u64)u1 << 32) | u0);
  u32 q1 = (u32)(q >> 32) + 1, q0 = (u32)q;
  u32 r = u0 - q1 * d;
  if (r > q0) { q1--; r += d; }
  if (r >= d) { q1++; r -= d; }
  *rem = r;
  return q1;
}

// rem[ND] = u[NU] mod d[ND] for d != 0 with dynamic size: Knuth D on a left-normalised copy of d.
// The dividend sits in a shift register T (NU+1 limbs, u << s bits): every step divides the top ND+1 limbs and moves
// T up by one limb, so all register indices are static while the loop itself stays ROLLED -- its trip count is the
// true number of quotient limbs (NU - ND + 1 + leading zero limbs of d), and the body exists once per call site
// (~130 SASS instructions instead of ~1.5 k unrolled: this function was half of the composition's code footprint).
// Returns false if d == 0.
template <int NU, int ND>
QF_HD bool fw_mod(u32 *rem, const u32 *u_in, const u32 *d_in) {
  static_assert(NU >= ND, "fw_mod");
  u32 d[ND];
  fw_copy<ND>(d, d_in);
  int lz = fw_lead_zero_limbs<ND>(d);
  if (lz == ND) return false;
  fw_shl_limbs<ND>(d, lz);
  int s = clz32(d[ND - 1]);
  fw_shl_bits<ND>(d, d, s);
  // d' = d << (32 lz + s) is normalised; u' = u << (32 lz + s).  T = u << s, read at the position where the first
  // window u'[jmax .. jmax+ND], jmax = NU + lz - ND, is the top of the register.
  u32 T[NU + 1];
  {
    u32 t[NU + 1];
    QF_UNROLL for (int i = 0; i < NU + 1; i++) t[i] = i < NU ? u_in[i] : 0;
    fw_shl_bits<NU + 1>(T, t, s);
  }
  const u32 dt = d[ND - 1], v = recip32(dt);
  const int cnt = NU - ND + 1 + lz;
  constexpr int B = NU - ND;  // window = T[B .. NU]
  QF_NOUNROLL for (int it = 0; it < cnt; it++) {
    // quotient limb from the two top limbs; top <= dt is an invariant
    u32 u1 = T[NU], u0 = T[NU - 1], r;
    u32 qh = (u1 >= dt) ? 0xffffffffu : div_2by1(u1, u0, dt, v, &r);
    if (qh != 0) {
      u32 bo = fw_submul1<ND>(T + B, d, qh);
      u32 top = T[NU];
      T[NU] = top - bo;
      if (top < bo) {
        // went negative: add back (at most twice)
        u32 c = fw_add<ND>(T + B, T + B, d);
        T[NU] += c;
        if (T[NU] != 0) { c = fw_add<ND>(T + B, T + B, d); T[NU] += c; }
      }
    }
    // the top limb is zero now: move the register up by one limb
    QF_UNROLL for (int i = NU; i > 0; i--) T[i] = T[i - 1];
    T[0] = 0;
  }
  // remainder' = T[B+1 .. NU] (after the last move); remainder = remainder' >> (32*lz + s)
  u32 r1[ND], r2[ND];
  QF_UNROLL for (int i = 0; i < ND; i++) r1[i] = T[B + 1 + i];
  fw_shr_bits<ND>(r2, r1, s);
  fw_shr_limbs<ND>(r2, lz);
  fw_copy<ND>(rem, r2);
  return true;
}

// r = a mod w, single limb divisor (w != 0)
template <int N>
QF_HD u32 fw_mod1(const u32 *a, u32 w) {
  int s = clz32(w);
  u32 d = w << s, v = recip32(d);
  // process a << s
  u32 r = s ? (a[N - 1] >> (32 - s)) : 0;
  // r < d holds since r < 2^s <= d... (d >= 2^31 >= r)
  QF_UNROLL for (int i = N - 1; i >= 0; i--) {
    u32 lo = s ? ((a[i] << s) | (i ? a[i - 1] >> (32 - s) : 0)) : a[i];
    u32 rem;
    div_2by1(r, lo, d, v, &rem);
    r = rem;
  }
  return r >> s;
}
// q[N] = a[N] / w exactly (w != 0, w | a)
template <int N>
QF_HD void fw_divexact1(u32 *q, const u32 *a, u32 w) {
  int t = ctz32(w);
  u32 wo = w >> t, wi = inv32(wo);
  u32 x[N];
  fw_shr_bits<N>(x, a, t);
  u32 bor = 0;
  QF_UNROLL for (int i = 0; i < N; i++) {
    u32 cur = x[i] - bor;
    u32 b1 = x[i] < bor ? 1u : 0u;
    u32 qi = cur * wi;
    q[i] = qi;
    bor = (u32)(((u64)qi * wo) >> 32) + b1;
  }
}

}  // namespace cofhe
