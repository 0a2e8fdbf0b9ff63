// cofhe_b200 -- per-element bodies of the batched entry points (generic path).  One call = one element of a
// batch laid out as the C ABI describes (include/cofhe_b200.h "Layout").  The CUDA kernels (kernels.cu) run one
// body per thread; tests/hostsim.cpp runs the same bodies in a CPU loop for the logic tests.
#pragma once
#include "qform.cuh"

namespace cofhe {

// shape of one form in a buffer: a[na] b[na] c[nc], limb-major with `n` elements per limb row
struct Shape {
  int na, nc;
};
QF_HD int form_limbs(const Shape &s) { return 2 * s.na + s.nc; }

template <int CAP>
QF_HDN void form_load(Err &e, Form<CAP> &f, const u32 *p, const Shape &s, size_t n, size_t idx) {
  z_load(e, f.a, p + idx, s.na, n, false);
  z_load(e, f.b, p + (size_t)s.na * n + idx, s.na, n, true);
  z_load(e, f.c, p + (size_t)2 * s.na * n + idx, s.nc, n, false);
}
template <int CAP>
QF_HDN void form_store(Err &e, const Form<CAP> &f, u32 *p, const Shape &s, size_t n, size_t idx) {
  z_store(e, f.a, p + idx, s.na, n, false);
  z_store(e, f.b, p + (size_t)s.na * n + idx, s.na, n, true);
  z_store(e, f.c, p + (size_t)2 * s.na * n + idx, s.nc, n, false);
}

// a and b rows only; c = (b^2 + |Delta|) / (4a) is recomputed (operands of the host-pointer add: include/cofhe_b200.h)
template <int CAP>
QF_HDN void form_load_ab(Err &e, const Group<CAP> &G, Form<CAP> &f, const u32 *p, const Shape &s, size_t n, size_t idx) {
  z_load(e, f.a, p + idx, s.na, n, false);
  z_load(e, f.b, p + (size_t)s.na * n + idx, s.na, n, true);
  Z<CAP> t, u;
  z_mul(e, t, f.b, f.b);
  z_add(e, u, t, G.absD);
  z_shr(e, t, u, 2);
  z_divexact<CAP>(e, f.c, t, f.a);
}

// out[idx] = in1[idx] o in2[idx]^(neg2 ? -1 : 1); no_c: the operands' c rows are not read
template <int CAP>
QF_HDN int op_compose(const Group<CAP> &G, const Shape &s, size_t n, size_t idx, const u32 *in1, const u32 *in2, u32 *out,
                      bool neg2, bool no_c = false) {
  Err e; e.code = 0;
  Form<CAP> f1, f2;
  if (no_c) {
    form_load_ab(e, G, f1, in1, s, n, idx);
    form_load_ab(e, G, f2, in2, s, n, idx);
  } else {
    form_load(e, f1, in1, s, n, idx);
    form_load(e, f2, in2, s, n, idx);
  }
  qf_compose(e, G, f1, f1, f2, neg2, false);
  form_store(e, f1, out, s, n, idx);
  return e.code;
}
template <int CAP>
QF_HDN int op_square(const Group<CAP> &G, const Shape &s, size_t n, size_t idx, const u32 *in, u32 *out) {
  Err e; e.code = 0;
  Form<CAP> f;
  form_load(e, f, in, s, n, idx);
  qf_compose(e, G, f, f, f, false, true);
  form_store(e, f, out, s, n, idx);
  return e.code;
}
template <int CAP>
QF_HDN int op_reduce(const Group<CAP> &G, const Shape &sin, const Shape &sout, size_t n, size_t idx, const u32 *in,
                     u32 *out) {
  (void)G;
  Err e; e.code = 0;
  Form<CAP> f;
  form_load(e, f, in, sin, n, idx);
  qf_reduce(e, f);
  form_store(e, f, out, sout, n, idx);
  return e.code;
}
// out[idx] = in[idx]^exp; exponent magnitude has el limbs; per_elem: exp is [el][n] (limb-major), sign in
// expneg[idx] (may be null); else one shared exponent exp[el] with sign shared_neg
template <int CAP>
QF_HDN int op_pow(const Group<CAP> &G, const Shape &s, size_t n, size_t idx, const u32 *in, u32 *out, const u32 *exp,
                  int el, bool per_elem, const unsigned char *expneg, bool shared_neg) {
  Err e; e.code = 0;
  Form<CAP> f, r;
  form_load(e, f, in, s, n, idx);
  if (per_elem) qf_pow(e, G, r, f, exp + idx, el, n, expneg ? expneg[idx] != 0 : false);
  else qf_pow(e, G, r, f, exp, el, 1, shared_neg);
  form_store(e, r, out, s, n, idx);
  return e.code;
}

}  // namespace cofhe
