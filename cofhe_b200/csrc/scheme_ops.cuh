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

}  // namespace cofhe
