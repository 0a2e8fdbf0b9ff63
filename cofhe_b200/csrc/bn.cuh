// cofhe_b200 -- generic (dynamic-length) multi-precision integers for device code.
//
// This is the ROBUST path: sign-magnitude integers of up to CAP 32-bit limbs held in thread-local memory,
// plain schoolbook loops, Knuth long division.  It backs every form operation for operand shapes the
// fixed-width register kernels (qf_fast.cuh) do not cover (unreduced inputs, tiny `a` / huge `c`, quotients
// wider than one limb, large discriminant tiers) and all the non-hot scheme arithmetic.  It replaces what
// the reference gets from libgmp's mpz_* (thirdparty/bicycl/src/bicycl/gmp_extras.inl:715-860).
//
// The code is plain C++ so the very same source also compiles with g++ for the CPU-side logic tests
// (tests/hostsim.cpp) -- that build is test infrastructure, never a product fallback.
#pragma once
#include <stdint.h>

#if defined(__CUDACC__)
#define QF_HD __host__ __device__ __forceinline__
#define QF_HDN __host__ __device__ __noinline__
#else
#define QF_HD inline
#define QF_HDN
#endif

namespace cofhe {

typedef uint32_t u32;
typedef uint64_t u64;
typedef int64_t i64;

// Compact variant of CL_HSM2k (reference CL_HSM2k.inl:255 `Cl_G`): the c1 half of a ciphertext -- form block 0 of a
// ciphertext batch -- lives in Cl(DeltaK), the c2 half in Cl(Delta).  A launch over a ciphertext batch of a compact context
// carries bits(|DeltaK|) in the high half of its `dbits` argument and the group of DeltaK right behind the group of
// Delta in device memory; every other launch passes a plain `dbits` (high half zero) and the group it means.
// Returns true when form block `y` is served by the second group (the caller then advances its group pointer by one).
QF_HD bool gs_pick(int &dbits, unsigned y) {
  const int k = dbits >> 16;
  dbits &= 0xffff;
  if (k && y == 0) { dbits = k; return true; }
  return false;
}

QF_HD int clz32(u32 x) {
#if defined(__CUDA_ARCH__)
  return __clz((int)x);
#else
  return x ? __builtin_clz(x) : 32;
#endif
}
QF_HD int clz64(u64 x) {
#if defined(__CUDA_ARCH__)
  return __clzll((long long)x);
#else
  return x ? __builtin_clzll(x) : 64;
#endif
}
QF_HD int ctz32(u32 x) {
#if defined(__CUDA_ARCH__)
  return x ? (__ffs((int)x) - 1) : 32;
#else
  return x ? __builtin_ctz(x) : 32;
#endif
}

// ------------------------------------------------------------------------------------------------
// magnitude (mpn-like) primitives on (pointer, length); lengths are "normalised" (top limb != 0, 0 = zero)
// ------------------------------------------------------------------------------------------------
QF_HD int bn_norm(const u32 *a, int n) {
  while (n > 0 && a[n - 1] == 0) n--;
  return n;
}
QF_HD int bn_cmp(const u32 *a, int na, const u32 *b, int nb) {
  if (na != nb) return na < nb ? -1 : 1;
  for (int i = na - 1; i >= 0; i--)
    if (a[i] != b[i]) return a[i] < b[i] ? -1 : 1;
  return 0;
}
QF_HD int bn_bits(const u32 *a, int na) { return na ? 32 * na - clz32(a[na - 1]) : 0; }
// r = a + b; r may alias a or b; r must hold max(na,nb)+1 limbs
QF_HD int bn_add(u32 *r, const u32 *a, int na, const u32 *b, int nb) {
  if (na < nb) { const u32 *t = a; a = b; b = t; int k = na; na = nb; nb = k; }
  u64 c = 0;
  int i = 0;
  for (; i < nb; i++) { c += (u64)a[i] + b[i]; r[i] = (u32)c; c >>= 32; }
  for (; i < na; i++) { c += a[i]; r[i] = (u32)c; c >>= 32; }
  if (c) r[i++] = (u32)c;
  return i;
}
// r = a - b, requires a >= b; r may alias a or b
QF_HD int bn_sub(u32 *r, const u32 *a, int na, const u32 *b, int nb) {
  i64 c = 0;
  int i = 0;
  for (; i < nb; i++) { c += (i64)a[i] - (i64)b[i]; r[i] = (u32)c; c >>= 32; }
  for (; i < na; i++) { c += (i64)a[i]; r[i] = (u32)c; c >>= 32; }
  return bn_norm(r, na);
}
// r = a * b (no aliasing); r holds na+nb limbs
QF_HD int bn_mul(u32 *r, const u32 *a, int na, const u32 *b, int nb) {
  if (na == 0 || nb == 0) return 0;
  for (int i = 0; i < na + nb; i++) r[i] = 0;
  for (int i = 0; i < na; i++) {
    u64 c = 0, ai = a[i];
    for (int j = 0; j < nb; j++) { c += ai * b[j] + r[i + j]; r[i + j] = (u32)c; c >>= 32; }
    r[i + nb] = (u32)c;
  }
  return bn_norm(r, na + nb);
}
// r = a << bits (r may alias a); r holds na + bits/32 + 1 limbs
QF_HD int bn_shl(u32 *r, const u32 *a, int na, unsigned bits) {
  if (na == 0) return 0;
  int ws = bits >> 5, bs = bits & 31;
  if (bs == 0) {
    for (int i = na - 1; i >= 0; i--) r[i + ws] = a[i];
  } else {
    r[na + ws] = a[na - 1] >> (32 - bs);
    for (int i = na - 1; i > 0; i--) r[i + ws] = (a[i] << bs) | (a[i - 1] >> (32 - bs));
    r[ws] = a[0] << bs;
  }
  for (int i = 0; i < ws; i++) r[i] = 0;
  return bn_norm(r, na + ws + (bs ? 1 : 0));
}
// r = a >> bits (floor on magnitude; r may alias a)
QF_HD int bn_shr(u32 *r, const u32 *a, int na, unsigned bits) {
  int ws = bits >> 5, bs = bits & 31;
  if (ws >= na) return 0;
  int n = na - ws;
  if (bs == 0) {
    for (int i = 0; i < n; i++) r[i] = a[i + ws];
  } else {
    for (int i = 0; i < n - 1; i++) r[i] = (a[i + ws] >> bs) | (a[i + ws + 1] << (32 - bs));
    r[n - 1] = a[na - 1] >> bs;
  }
  return bn_norm(r, n);
}
// number of trailing zero bits (a != 0)
QF_HD int bn_ctz(const u32 *a, int na) {
  for (int i = 0; i < na; i++)
    if (a[i]) return 32 * i + ctz32(a[i]);
  return 0;
}

// Knuth algorithm D.  q gets na-nb+1 limbs, r gets nb limbs; `an` is scratch of na+1 limbs, `bs` of nb limbs.
// q and r must not alias a or b.  Returns lengths through nq/nr.  Single exit on purpose (the 64-bit divisions are
// calls on the device; see DESIGN.md "device-code rules").
inline QF_HDN void bn_divrem(u32 *q, int *nq, u32 *r, int *nr, const u32 *a, int na, const u32 *b, int nb, u32 *an,
                      u32 *bs) {
  if (bn_cmp(a, na, b, nb) < 0) {
    for (int i = 0; i < na; i++) r[i] = a[i];
    *nr = na;
    *nq = 0;
  } else if (nb == 1) {
    u64 rem = 0;
    u32 d = b[0];
    for (int i = na - 1; i >= 0; i--) {
      u64 cur = (rem << 32) | a[i];
      q[i] = (u32)(cur / d);
      rem = cur % d;
    }
    r[0] = (u32)rem;
    *nr = rem ? 1 : 0;
    *nq = bn_norm(q, na);
  } else {
  int s = clz32(b[nb - 1]);
  if (s) {
    for (int i = nb - 1; i > 0; i--) bs[i] = (b[i] << s) | (b[i - 1] >> (32 - s));
    bs[0] = b[0] << s;
    an[na] = a[na - 1] >> (32 - s);
    for (int i = na - 1; i > 0; i--) an[i] = (a[i] << s) | (a[i - 1] >> (32 - s));
    an[0] = a[0] << s;
  } else {
    for (int i = 0; i < nb; i++) bs[i] = b[i];
    for (int i = 0; i < na; i++) an[i] = a[i];
    an[na] = 0;
  }
  u64 bt = bs[nb - 1], bt2 = bs[nb - 2];
  for (int j = na - nb; j >= 0; j--) {
    u64 num = ((u64)an[j + nb] << 32) | an[j + nb - 1];
    u64 qhat = num / bt, rhat = num % bt;
    while (qhat >= (1ull << 32) || qhat * bt2 > ((rhat << 32) | an[j + nb - 2])) {
      qhat--;
      rhat += bt;
      if (rhat >= (1ull << 32)) break;
    }
    // multiply and subtract
    i64 borrow = 0;
    u64 carry = 0;
    for (int i = 0; i < nb; i++) {
      u64 p = qhat * bs[i] + carry;
      carry = p >> 32;
      i64 t = (i64)an[i + j] - (i64)(u32)p + borrow;
      an[i + j] = (u32)t;
      borrow = t >> 32;
    }
    i64 t = (i64)an[j + nb] - (i64)carry + borrow;
    an[j + nb] = (u32)t;
    if (t < 0) {
      qhat--;
      u64 c = 0;
      for (int i = 0; i < nb; i++) { c += (u64)an[i + j] + bs[i]; an[i + j] = (u32)c; c >>= 32; }
      an[j + nb] += (u32)c;
    }
    q[j] = (u32)qhat;
  }
  *nq = bn_norm(q, na - nb + 1);
  if (s) {
    for (int i = 0; i < nb - 1; i++) r[i] = (an[i] >> s) | (an[i + 1] << (32 - s));
    r[nb - 1] = an[nb - 1] >> s;
  } else {
    for (int i = 0; i < nb; i++) r[i] = an[i];
  }
  *nr = bn_norm(r, nb);
  }
}

// ------------------------------------------------------------------------------------------------
// signed integers, capacity CAP limbs.  `err` (sticky, in Err) is raised when a result would not fit.
// ------------------------------------------------------------------------------------------------
struct Err {
  int code;  // 0 ok; 1 capacity overflow; 2 inexact division; 3 domain
};

template <int CAP>
struct Z {
  int n;    // limbs in use (0 = zero)
  int neg;  // 1 if negative (never set for zero)
  u32 d[CAP];
};

template <int CAP>
QF_HD void z_zero(Z<CAP> &r) { r.n = 0; r.neg = 0; }
template <int CAP>
QF_HD void z_set_u32(Z<CAP> &r, u32 v) { r.d[0] = v; r.n = v ? 1 : 0; r.neg = 0; }
template <int CAP>
QF_HD void z_set_u64(Z<CAP> &r, u64 v) {
  r.d[0] = (u32)v; r.d[1] = (u32)(v >> 32);
  r.n = (v >> 32) ? 2 : (v ? 1 : 0); r.neg = 0;
}
template <int CA, int CB>
QF_HDN void z_copy(Err &e, Z<CA> &r, const Z<CB> &a) {
  if (a.n > CA) { e.code = 1; r.n = 0; r.neg = 0; return; }
  for (int i = 0; i < a.n; i++) r.d[i] = a.d[i];
  r.n = a.n; r.neg = a.neg;
}
template <int CAP>
QF_HD void z_neg(Z<CAP> &r) { if (r.n) r.neg ^= 1; }
template <int CAP>
QF_HD int z_sgn(const Z<CAP> &a) { return a.n ? (a.neg ? -1 : 1) : 0; }
template <int CAP>
QF_HD int z_bits(const Z<CAP> &a) { return bn_bits(a.d, a.n); }
template <int CAP>
QF_HD bool z_is_one(const Z<CAP> &a) { return a.n == 1 && a.d[0] == 1 && !a.neg; }
template <int CAP>
QF_HD bool z_odd(const Z<CAP> &a) { return a.n && (a.d[0] & 1); }
template <int CA, int CB>
QF_HD int z_cmpabs(const Z<CA> &a, const Z<CB> &b) { return bn_cmp(a.d, a.n, b.d, b.n); }
template <int CA, int CB>
QF_HDN int z_cmp(const Z<CA> &a, const Z<CB> &b) {
  int sa = z_sgn(a), sb = z_sgn(b);
  if (sa != sb) return sa < sb ? -1 : 1;
  int c = bn_cmp(a.d, a.n, b.d, b.n);
  return sa < 0 ? -c : c;
}

// r = a + (bneg ? -b : b) with explicit sign of b; r may alias a or b
template <int CAP, int CA, int CB>
QF_HDN void z_addsub(Err &e, Z<CAP> &r, const Z<CA> &a, const Z<CB> &b, int flip) {
  int bneg = b.n ? (b.neg ^ flip) : 0;
  int mx = a.n > b.n ? a.n : b.n;
  if (a.neg == bneg || a.n == 0 || b.n == 0) {
    if (mx + 1 > CAP) {
      // may still fit: check by computing the carry cheaply -- be conservative only when really needed
      u64 c = 0;
      for (int i = 0; i < mx; i++) { c += (u64)(i < a.n ? a.d[i] : 0) + (i < b.n ? b.d[i] : 0); c >>= 32; }
      if (mx > CAP || c) { e.code = 1; r.n = 0; r.neg = 0; return; }
      u64 cc = 0;
      for (int i = 0; i < mx; i++) {
        cc += (u64)(i < a.n ? a.d[i] : 0) + (i < b.n ? b.d[i] : 0);
        r.d[i] = (u32)cc; cc >>= 32;
      }
      r.neg = a.n ? a.neg : bneg;
      r.n = bn_norm(r.d, mx);
      if (!r.n) r.neg = 0;
      return;
    }
    int sg = a.n ? a.neg : bneg;
    r.n = bn_add(r.d, a.d, a.n, b.d, b.n);
    r.neg = r.n ? sg : 0;
    return;
  }
  if (mx > CAP) { e.code = 1; r.n = 0; r.neg = 0; return; }
  int c = bn_cmp(a.d, a.n, b.d, b.n);
  if (c == 0) { r.n = 0; r.neg = 0; return; }
  if (c > 0) { int sg = a.neg; r.n = bn_sub(r.d, a.d, a.n, b.d, b.n); r.neg = sg; }
  else { int sg = bneg; r.n = bn_sub(r.d, b.d, b.n, a.d, a.n); r.neg = sg; }
}
template <int CAP, int CA, int CB>
QF_HD void z_add(Err &e, Z<CAP> &r, const Z<CA> &a, const Z<CB> &b) { z_addsub(e, r, a, b, 0); }
template <int CAP, int CA, int CB>
QF_HD void z_sub(Err &e, Z<CAP> &r, const Z<CA> &a, const Z<CB> &b) { z_addsub(e, r, a, b, 1); }

// r = a * b; r must not alias a or b
template <int CAP, int CA, int CB>
QF_HDN void z_mul(Err &e, Z<CAP> &r, const Z<CA> &a, const Z<CB> &b) {
  if (a.n == 0 || b.n == 0) { r.n = 0; r.neg = 0; return; }
  if (a.n + b.n > CAP) {
    // the product may still fit in CAP limbs (top limb zero): compute into a bounded loop
    int bits = bn_bits(a.d, a.n) + bn_bits(b.d, b.n);
    if (bits - 1 > 32 * CAP || a.n + b.n > CAP + 1) { e.code = 1; r.n = 0; r.neg = 0; return; }
    for (int i = 0; i < CAP; i++) r.d[i] = 0;
    u32 top = 0;
    for (int i = 0; i < a.n; i++) {
      u64 c = 0, ai = a.d[i];
      for (int j = 0; j < b.n; j++) {
        if (i + j < CAP) { c += ai * b.d[j] + r.d[i + j]; r.d[i + j] = (u32)c; c >>= 32; }
        else { c += ai * b.d[j]; top |= (u32)c; c >>= 32; }
      }
      if (i + b.n < CAP) r.d[i + b.n] = (u32)c; else top |= (u32)c;
    }
    if (top) { e.code = 1; r.n = 0; r.neg = 0; return; }
    r.n = bn_norm(r.d, CAP);
    r.neg = a.neg ^ b.neg;
    return;
  }
  r.n = bn_mul(r.d, a.d, a.n, b.d, b.n);
  r.neg = a.neg ^ b.neg;
}
template <int CAP, int CA>
QF_HDN void z_mul_u32(Err &e, Z<CAP> &r, const Z<CA> &a, u32 w) {  // r may alias a
  if (a.n == 0 || w == 0) { r.n = 0; r.neg = 0; return; }
  u64 c = 0;
  int n = a.n, sg = a.neg;
  if (n > CAP) { e.code = 1; r.n = 0; r.neg = 0; return; }
  for (int i = 0; i < n; i++) { c += (u64)a.d[i] * w; r.d[i] = (u32)c; c >>= 32; }
  if (c) { if (n >= CAP) { e.code = 1; r.n = 0; r.neg = 0; return; } r.d[n++] = (u32)c; }
  r.n = n; r.neg = sg;
}
template <int CAP, int CA>
QF_HDN void z_shl(Err &e, Z<CAP> &r, const Z<CA> &a, unsigned bits) {  // r may alias a
  if (a.n == 0) { r.n = 0; r.neg = 0; return; }
  if (bn_bits(a.d, a.n) + (int)bits > 32 * CAP) { e.code = 1; r.n = 0; r.neg = 0; return; }
  int sg = a.neg;
  int ws = bits >> 5, bs = bits & 31, na = a.n;
  // like bn_shl but never touches index >= CAP
  if (bs == 0) {
    for (int i = na - 1; i >= 0; i--) r.d[i + ws] = a.d[i];
  } else {
    u32 hi = a.d[na - 1] >> (32 - bs);
    if (hi) r.d[na + ws] = hi;
    for (int i = na - 1; i > 0; i--) r.d[i + ws] = (a.d[i] << bs) | (a.d[i - 1] >> (32 - bs));
    r.d[ws] = a.d[0] << bs;
    if (hi) na++;
  }
  for (int i = 0; i < ws; i++) r.d[i] = 0;
  r.n = bn_norm(r.d, na + ws);
  r.neg = sg;
}
// magnitude shift right (truncation toward zero); exact when the dropped bits are zero
template <int CAP, int CA>
QF_HDN void z_shr(Err &e, Z<CAP> &r, const Z<CA> &a, unsigned bits) {
  int sg = a.neg;
  if (a.n - (int)(bits >> 5) > CAP) { e.code = 1; r.n = 0; r.neg = 0; return; }
  r.n = bn_shr(r.d, a.d, a.n, bits);
  r.neg = r.n ? sg : 0;
}

// truncated division: a = q*b + r, |r| < |b|, sign(r) = sign(a).  q, r must not alias a, b.
template <int CAP, int CQ, int CR, int CA, int CB>
QF_HDN void z_tdiv_qr(Err &e, Z<CQ> &q, Z<CR> &r, const Z<CA> &a, const Z<CB> &b) {
  // NB single exit on purpose: see DESIGN.md "device-code rules" (early returns around calls under divergence)
  int mode;  // 0: error, 1: |a| < |b|, 2: divide
  if (b.n == 0) { e.code = 3; mode = 0; }
  else if (a.n > CAP || b.n > CAP) { e.code = 1; mode = 0; }
  else if (bn_cmp(a.d, a.n, b.d, b.n) < 0) {
    if (a.n > CR) { e.code = 1; mode = 0; } else mode = 1;
  } else if (a.n - b.n + 1 > CQ || b.n > CR) { e.code = 1; mode = 0; }
  else mode = 2;
  if (mode == 2) {
    u32 an[CAP + 1], bs[CAP];
    int nq, nr;
    bn_divrem(q.d, &nq, r.d, &nr, a.d, a.n, b.d, b.n, an, bs);
    q.n = nq; q.neg = nq ? (a.neg ^ b.neg) : 0;
    r.n = nr; r.neg = nr ? a.neg : 0;
  } else if (mode == 1) {
    for (int i = 0; i < a.n; i++) r.d[i] = a.d[i];
    r.n = a.n; r.neg = a.neg; q.n = 0; q.neg = 0;
  } else {
    q.n = r.n = 0; q.neg = r.neg = 0;
  }
}
// floor division: a = q*b + r with 0 <= r < b for b > 0
template <int CAP, int CQ, int CR, int CA, int CB>
QF_HDN void z_fdiv_qr(Err &e, Z<CQ> &q, Z<CR> &r, const Z<CA> &a, const Z<CB> &b) {
  z_tdiv_qr<CAP>(e, q, r, a, b);
  if (r.n && (r.neg != b.neg)) {
    // q -= 1 ; r += b
    Z<2> one; z_set_u32(one, 1);
    z_sub(e, q, q, one);
    z_add(e, r, r, b);
  }
}
template <int CAP, int CQ, int CA, int CB>
QF_HDN void z_divexact(Err &e, Z<CQ> &q, const Z<CA> &a, const Z<CB> &b) {
  Z<CAP> r;
  z_tdiv_qr<CAP>(e, q, r, a, b);
  if (r.n) e.code = 2;
}
// r = a mod b in [0, |b|)
template <int CAP, int CR, int CA, int CB>
QF_HDN void z_mod(Err &e, Z<CR> &r, const Z<CA> &a, const Z<CB> &b) {
  Z<CAP> q;
  z_tdiv_qr<CAP>(e, q, r, a, b);
  if (r.neg) { if (b.neg) z_sub(e, r, r, b); else z_add(e, r, r, b); }
}

// load/store against the fixed-width SoA layout of the C ABI (include/cofhe_b200.h "Layout"):
// limb i of element idx at p[i*stride + idx]; signed values in two's complement over nl limbs.
template <int CAP>
QF_HDN void z_load(Err &e, Z<CAP> &r, const u32 *p, int nl, size_t stride, bool twos) {
  bool neg = twos && (p[(size_t)(nl - 1) * stride] >> 31);
  int n = nl;
  if (!neg) {
    while (n > 0 && p[(size_t)(n - 1) * stride] == 0) n--;
    if (n > CAP) { e.code = 1; r.n = 0; r.neg = 0; return; }
    for (int i = 0; i < n; i++) r.d[i] = p[(size_t)i * stride];
    r.n = n; r.neg = 0;
    return;
  }
  while (n > 1 && p[(size_t)(n - 1) * stride] == 0xffffffffu) n--;
  if (n > CAP) { e.code = 1; r.n = 0; r.neg = 0; return; }
  u64 c = 1;
  for (int i = 0; i < n; i++) { c += (u32)~p[(size_t)i * stride]; r.d[i] = (u32)c; c >>= 32; }
  if (c && n < CAP) { r.d[n] = 1; n++; } else if (c) { e.code = 1; }
  r.n = bn_norm(r.d, n); r.neg = r.n ? 1 : 0;
}
template <int CAP>
QF_HDN void z_store(Err &e, const Z<CAP> &a, u32 *p, int nl, size_t stride, bool twos) {
  int bits = bn_bits(a.d, a.n);
  if (a.neg && !twos) e.code = 3;
  if (twos) {
    // representable range: [-2^(32nl-1), 2^(32nl-1))
    bool ok = bits < 32 * nl || (a.neg && bits == 32 * nl && bn_ctz(a.d, a.n) == 32 * nl - 1);
    if (!ok) e.code = 1;
  } else if (bits > 32 * nl) e.code = 1;
  if (!a.neg) {
    for (int i = 0; i < nl; i++) p[(size_t)i * stride] = i < a.n ? a.d[i] : 0;
  } else {
    u64 c = 1;
    for (int i = 0; i < nl; i++) { c += (u32)~(i < a.n ? a.d[i] : 0); p[(size_t)i * stride] = (u32)c; c >>= 32; }
  }
}

}  // namespace cofhe
