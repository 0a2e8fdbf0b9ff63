// cofhe_b200 -- the float <-> plaintext codec of CPUCryptoSystem (reference include/x86_64/cpu_cryptosystem.inl:49-87:
// map_to_positive / map_back, behind make_plaintext / get_float_from_plaintext :114-122) as per-element integer code.
//
// The reference computes in GMP's mpf with the DEFAULT precision (mpf_init: 64 bits -> _mp_prec = 2 limbs, every
// operation keeps prec + 1 = 3 limbs of 64 bits) and scaling_factor = 2^0 (cpu_cryptosystem.hpp:158), M = 2^k:
//   make_plaintext(x)   = trunc(x)                           x >= 0     (mpz_set_f truncates towards zero)
//                       = trunc(M (-) |x|)                    x <  0     where (-) is mpf_sub's subtraction of magnitudes:
//        with uexp = k/64 + 1, vexp = 64-bit limbs of the integer part of |x| (<= 0 below 1) and ediff = uexp - vexp,
//        ediff <= 1: exact;  ediff >= 3: |x| is dropped entirely (result M);  ediff == 2: only the TOP limb of |x| takes part
//        (mpf/sub.c: "If V extends beyond PREC, ignore the part that does") -- at k = 128 a negative x maps to M + trunc(x),
//        at k >= 192 every negative |x| < 2^64 maps to M.  That is what the reference encrypts, so it is reproduced.
//   get_float(z) = (float) RZ_double(v),  v = z' (z' < M/2) or z' - M, z' = z with all but its top three 64-bit limbs cleared
//        (mpf_set_z), RZ_double = truncation to 53 bits (mpf_get_d), the cast to float rounds to nearest even.
// Values are sign-magnitude: magnitude limbs (little endian u32) plus a sign flag, as the plaintext batches of the wire
// kernels.  The same functions compile for the host test build (tests/hostsim).
#pragma once
#include "bn.cuh"

namespace cofhe {

// mag[0..nl) = |result|, returns the sign (1: negative).  x must be finite.
QF_HD int codec_make_plaintext(float x, int k, u32 *mag, int nl) {
  for (int i = 0; i < nl; i++) mag[i] = 0;
  u32 bits;
#if defined(__CUDA_ARCH__)
  bits = __float_as_uint(x);
#else
  { union { float f; u32 u; } cv; cv.f = x; bits = cv.u; }
#endif
  const bool neg = (bits >> 31) != 0;
  const int be = (int)((bits >> 23) & 0xff);
  u32 man = bits & 0x7fffffu;
  if (be == 0 && man == 0) return 0;  // +-0
  int e;                              // |x| = man * 2^e
  if (be == 0) e = -149; else { man |= 0x800000u; e = be - 150; }
  // integer part ip = floor(|x|) as up to 128 bits, frac: is there a fractional part
  u32 ip[5] = {0, 0, 0, 0, 0};
  bool frac = false;
  if (e >= 0) {
    int w = e >> 5, s = e & 31;
    ip[w] = man << s;
    if (s) ip[w + 1] = man >> (32 - s);
  } else if (e > -24) {
    ip[0] = man >> (-e);
    frac = (man & ((1u << (-e)) - 1u)) != 0;
  } else frac = true;
  int ipbits = 0;
  for (int i = 0; i < 5; i++) if (ip[i]) ipbits = 32 * i + 32 - clz32(ip[i]);
  if (!neg) {
    for (int i = 0; i < nl && i < 5; i++) mag[i] = ip[i];
    return 0;
  }
  // ---- x < 0: M (-) |x| as mpf_sub does it
  const int uexp = k / 64 + 1;
  // exponent of |x| in 64-bit limbs, as mpf normalises it: |x| in [2^(64(vexp-1)), 2^(64 vexp)); <= 0 below 1
  int vexp_real;
  if (ipbits) vexp_real = (ipbits + 63) / 64;
  else {
    int tb = e + (32 - clz32(man));  // |x| in [2^(tb-1), 2^tb), tb <= 0
    vexp_real = -((-tb) / 64);
  }
  const int ediff = uexp - vexp_real;
  // v: the part of |x| that takes part, as (vi: integer part, vf: non-zero fraction)
  u32 vi[5] = {ip[0], ip[1], ip[2], ip[3], ip[4]};
  bool vf = frac;
  bool drop_all = false;
  if (ediff >= 3) drop_all = true;
  else if (ediff == 2) {
    // only the top 64-bit limb of V (the mpf of a double has two limbs: the top one holds its leading bits)
    vf = false;
    if (vexp_real >= 1) {
      int low = 64 * (vexp_real - 1);  // bits below the top limb of the integer part
      for (int i = 0; i < 5; i++) {
        int lo = 32 * i;
        if (lo + 32 <= low) vi[i] = 0;
        else if (lo < low) vi[i] &= ~((1u << (low - lo)) - 1u);
      }
    } else {
      // |x| < 1 and its top limb is a fraction limb: V' is a non-zero fraction
      for (int i = 0; i < 5; i++) vi[i] = 0;
      vf = true;
    }
  }
  // R = M - V (exact); result = trunc(R) towards zero.  M = 2^k.
  // c = vi + (vf ? 1 : 0) = ceil(V);  R >= 0: M - c ; R < 0 (V > M): -(floor(V) - M)
  u32 m[6] = {0, 0, 0, 0, 0, 0};  // M as limbs when k < 160, else handled separately
  if (drop_all) {
    if (k / 32 < nl) mag[k / 32] = 1u << (k & 31);
    return 0;
  }
  bool v_gt_m = false;
  if (k < 160) {
    m[k / 32] = 1u << (k & 31);
    // compare floor(V) with M
    int c = 0;
    for (int i = 4; i >= 0 && c == 0; i--) c = vi[i] > m[i] ? 1 : (vi[i] < m[i] ? -1 : 0);
    v_gt_m = c > 0 || (c == 0 && vf);
    if (c == 0 && !vf) return 0;  // M - M = 0
    if (v_gt_m) {
      // -(floor(V) - M)
      u64 br = 0;
      for (int i = 0; i < 5; i++) { u64 d = (u64)vi[i] - m[i] - br; if (i < nl) mag[i] = (u32)d; br = (d >> 63) & 1; }
      bool z = true;
      for (int i = 0; i < nl; i++) z = z && mag[i] == 0;
      return z ? 0 : 1;
    }
  }
  // M - ceil(V), M = 2^k with k possibly beyond the 160 bits of vi: subtract from the (nl-limb) image of 2^k
  u32 cv[6] = {vi[0], vi[1], vi[2], vi[3], vi[4], 0};
  if (vf) { u64 cy = 1; for (int i = 0; i < 6; i++) { u64 t = (u64)cv[i] + cy; cv[i] = (u32)t; cy = t >> 32; } }
  u64 br = 0;
  for (int i = 0; i < nl; i++) {
    u32 mi = (i == k / 32) ? (1u << (k & 31)) : 0u;
    u64 d = (u64)mi - (i < 6 ? cv[i] : 0u) - br;
    mag[i] = (u32)d;
    br = (d >> 63) & 1;
  }
  return 0;
}

// z = (sign ? -1 : 1) * mag[0..nl).  Returns get_float_from_plaintext(z).
QF_HD float codec_get_float(const u32 *mag_in, int nl, int sign, int k) {
  // work on a copy of at most 64 limbs (2048 bits)
  u32 z[66];
  int n = nl > 64 ? 64 : nl;
  for (int i = 0; i < n; i++) z[i] = mag_in[i];
  z[n] = 0; z[n + 1] = 0;
  while (n > 0 && z[n - 1] == 0) n--;
  if (n == 0) return 0.0f;
  // mpf_set_z keeps the top three 64-bit limbs
  int n64 = (n + 1) / 2;
  if (n64 > 3) for (int i = 0; i < 2 * (n64 - 3); i++) z[i] = 0;
  // v = z' (z' < M/2, or z negative: the reference compares the signed value) or z' - M
  bool vneg = sign != 0;
  if (!sign) {
    // z' >= 2^(k-1) ?
    bool ge = false;
    int hb = k - 1;
    for (int i = n - 1; i >= 0 && !ge; i--) {
      if (z[i] == 0) continue;
      int top = 32 * i + 31 - clz32(z[i]);
      ge = top >= hb;
      break;
    }
    if (ge) {
      // v = z' - M: negative when z' < M
      int mw = k / 32;
      u32 mb = 1u << (k & 31);
      int c = 0;  // compare z' with M
      for (int i = (n > mw + 1 ? n : mw + 1) - 1; i >= 0 && c == 0; i--) {
        u32 zi = i < n ? z[i] : 0u, mi = i == mw ? mb : 0u;
        c = zi > mi ? 1 : (zi < mi ? -1 : 0);
      }
      if (c == 0) return 0.0f;
      int top = n > mw + 1 ? n : mw + 1;
      if (top > 65) top = 65;
      u64 br = 0;
      for (int i = 0; i < top; i++) {
        u32 zi = i < n ? z[i] : 0u, mi = i == mw ? mb : 0u;
        u64 d = c > 0 ? (u64)zi - mi - br : (u64)mi - zi - br;
        z[i] = (u32)d;
        br = (d >> 63) & 1;
      }
      n = top;
      while (n > 0 && z[n - 1] == 0) n--;
      vneg = c < 0;
    }
  } else {
    // negative plaintext: mpf_cmp(num, M/2) < 0 holds, v = z
  }
  // top 53 bits, truncated
  int tb = 32 * (n - 1) + 31 - clz32(z[n - 1]);  // index of the top bit
  u64 m53 = 0;
  for (int b = 0; b < 53; b++) {
    int pos = tb - b;
    u32 bit = pos >= 0 ? (z[pos >> 5] >> (pos & 31)) & 1u : 0u;
    m53 = (m53 << 1) | bit;
  }
  int sh = tb - 52;  // value = m53 * 2^sh
  double d = (double)m53;
  // scale by 2^sh exactly (|sh| small enough for doubles here: tb < 2112)
  while (sh > 0) { int s = sh > 60 ? 60 : sh; d *= (double)((u64)1 << s); sh -= s; }
  while (sh < 0) { int s = -sh > 60 ? 60 : -sh; d /= (double)((u64)1 << s); sh += s; }
  float f = (float)d;
  return vneg ? -f : f;
}

}  // namespace cofhe
