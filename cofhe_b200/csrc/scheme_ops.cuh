// cofhe_b200 -- per-element bodies of the scheme-level entry points (one element of a batch per call; the CUDA
// kernels run one body per thread, tests/hostsim runs them in a CPU loop).
#pragma once
#include "ops.cuh"
#include "scheme.cuh"

namespace cofhe {

// plaintext / result integers modulo 2^k cross the ABI as [ml][n] limb rows (ml >= 1); missing high limbs are zero
template <int KL>
QF_HD void mk_load(u32 *r, const u32 *p, int ml, size_t n, size_t idx) {
  QF_UNROLL for (int i = 0; i < KL; i++) r[i] = i < ml ? p[(size_t)i * n + idx] : 0;
}
template <int KL>
QF_HD void mk_store(const u32 *r, u32 *p, int ml, size_t n, size_t idx) {
  QF_UNROLL for (int i = 0; i < KL; i++) if (i < ml) p[(size_t)i * n + idx] = r[i];
  for (int i = KL; i < ml; i++) p[(size_t)i * n + idx] = 0;
}

// out[idx] = reduce(f^m[idx])    (CL_HSM2k::power_of_f followed by the reduction every consumer applies)
template <int KL, int CAP>
QF_HDN int op_power_of_f(const Group<CAP> &G, const Scheme<CAP> &S, const Shape &s, size_t n, size_t idx, const u32 *m, int ml,
                         u32 *out) {
  Err e; e.code = 0;
  u32 mm[KL];
  mk_load<KL>(mm, m, ml, n, idx);
  Form<CAP> f;
  sch_power_of_f<KL, CAP>(e, G, S, f, mm);
  qf_reduce(e, f);
  form_store(e, f, out, s, n, idx);
  return e.code;
}

// out[idx] = dlog_F(in[idx]) in [0, 2^k)
template <int KL, int CAP>
QF_HDN int op_dlog(const Group<CAP> &G, const Scheme<CAP> &S, const Shape &s, size_t n, size_t idx, const u32 *in, u32 *out,
                   int ml) {
  Err e; e.code = 0;
  Form<CAP> f;
  form_load(e, f, in, s, n, idx);
  u32 m[KL];
  sch_dlog<KL, CAP>(e, G, S, m, f);
  mk_store<KL>(m, out, ml, n, idx);
  return e.code;
}

// ---- compact variant: Cl(DeltaK) -> Cl(Delta) (CL_HSM2k::from_Cl_DeltaK_to_Cl_Delta, CL_HSM2k.inl:458-474) -----------------
// Step 1: QFI::lift_2exp(k+1) (qfi.inl:263-271) = prime_to_2exp (:590-610: make a odd), b <- b 2^(k+1), c <- c 2^(2k+2),
// reduction; chi[idx] = 1 when a or c of the lifted reduced form is 3 mod 4 (CL_HSM2k.inl:461).
template <int CAP>
QF_HDN int op_lift_2exp(const Shape &s, size_t n, size_t idx, int k, const u32 *in, u32 *out, unsigned char *chi) {
  Err e; e.code = 0;
  Form<CAP> f;
  form_load(e, f, in, s, n, idx);
  if (!z_odd(f.a)) {
    Z<CAP> t;
    if (!z_odd(f.c)) {  // (a, b, c) -> (a + b + c, -b - 2a, a): first c <- a + b + c and b <- b + 2a, then the swap below
      z_add(e, t, f.c, f.a);
      z_add(e, f.c, t, f.b);
      z_add(e, t, f.b, f.a);
      z_add(e, f.b, t, f.a);
    }
    // (a, b, c) -> (c, -b, a)
    z_neg(f.b);
    z_copy(e, t, f.a);
    z_copy(e, f.a, f.c);
    z_copy(e, f.c, t);
  }
  z_shl(e, f.b, f.b, (unsigned)(k + 1));
  z_shl(e, f.c, f.c, (unsigned)(2 * k + 2));
  qf_reduce(e, f);
  chi[idx] = (unsigned char)(((f.a.n ? f.a.d[0] & 3u : 0u) == 3u || (f.c.n ? f.c.d[0] & 3u : 0u) == 3u) ? 1 : 0);
  form_store(e, f, out, s, n, idx);
  return e.code;
}
// Step 3 (step 2 = f^(2^k), k squarings in Cl(Delta): raise_to_power_M, CL_HSM2k.inl:641-648), for the elements whose
// chi is set: compose with g = (4, 4, 1 - DeltaK 2^(2(k-1)))  (CL_HSM2k.inl:464-473)
template <int CAP>
QF_HDN int op_lift_fix(const Group<CAP> &G, const Scheme<CAP> &S, const Shape &s, size_t n, size_t idx, u32 *io) {
  Err e; e.code = 0;
  Form<CAP> f, g;
  form_load(e, f, io, s, n, idx);
  z_set_u32(g.a, 4);
  z_set_u32(g.b, 4);
  Z<CAP> t, one;
  z_set_u32(one, 1);
  z_shl(e, t, S.absDK, (unsigned)(2 * (S.k - 1)));
  z_add(e, g.c, t, one);
  qf_compose(e, G, f, f, g, false, false);
  form_store(e, f, io, s, n, idx);
  return e.code;
}

}  // namespace cofhe
