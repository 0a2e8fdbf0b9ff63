// TEST INFRASTRUCTURE ONLY: compiles the device headers (cofhe_b200/csrc/*.cuh) with g++ and runs the
// per-element op bodies in a CPU loop so that the arithmetic LOGIC can be checked against the oracle in the
// GPU-less build container (`pytest -m "not gpu"`).  Nothing in the product path loads this library; the
// product (libcofhe_b200.so) has no CPU execution path at all.
#include <cstddef>
#include <cstring>
#include <vector>
#include "ops.cuh"
#include "hostsim_extra.inc"
#include "coop_sim.inc"
#include "codec.cuh"

using namespace cofhe;

template <int CAP>
static void mk_group(Group<CAP> &G, const u32 *absD, int nd) {
  G.absD.n = bn_norm(absD, nd);
  G.absD.neg = 0;
  for (int i = 0; i < G.absD.n; i++) G.absD.d[i] = absD[i];
  G.dbits = bn_bits(G.absD.d, G.absD.n);
}

#define DISPATCH(cap, ...)                \
  switch (cap) {                           \
    case 32: { constexpr int CAP = 32; __VA_ARGS__; } break;   \
    case 48: { constexpr int CAP = 48; __VA_ARGS__; } break;   \
    case 136: { constexpr int CAP = 136; __VA_ARGS__; } break; \
    case 216: { constexpr int CAP = 216; __VA_ARGS__; } break; \
    default: return -1;                    \
  }

extern "C" {

int hs_compose(int cap, const u32 *absD, int nd, int na, int nc, size_t n, const u32 *in1, const u32 *in2, u32 *out,
               int neg2) {
  int rc = 0;
  DISPATCH(cap, {
    Group<CAP> G; mk_group(G, absD, nd); Shape s{na, nc};
    for (size_t i = 0; i < n; i++) { int c = op_compose<CAP>(G, s, n, i, in1, in2, out, neg2 != 0); if (c && !rc) rc = c; }
  })
  return rc;
}
int hs_square(int cap, const u32 *absD, int nd, int na, int nc, size_t n, const u32 *in, u32 *out) {
  int rc = 0;
  DISPATCH(cap, {
    Group<CAP> G; mk_group(G, absD, nd); Shape s{na, nc};
    for (size_t i = 0; i < n; i++) { int c = op_square<CAP>(G, s, n, i, in, out); if (c && !rc) rc = c; }
  })
  return rc;
}
int hs_reduce(int cap, const u32 *absD, int nd, int na_in, int nc_in, int na, int nc, size_t n, const u32 *in, u32 *out) {
  int rc = 0;
  DISPATCH(cap, {
    Group<CAP> G; mk_group(G, absD, nd); Shape s{na, nc};
    Shape sin{na_in, nc_in};
    for (size_t i = 0; i < n; i++) { int c = op_reduce<CAP>(G, sin, s, n, i, in, out); if (c && !rc) rc = c; }
  })
  return rc;
}
int hs_pow(int cap, const u32 *absD, int nd, int na, int nc, size_t n, const u32 *in, u32 *out, const u32 *exp, int el,
           int per_elem, const unsigned char *expneg, int shared_neg) {
  int rc = 0;
  DISPATCH(cap, {
    Group<CAP> G; mk_group(G, absD, nd); Shape s{na, nc};
    for (size_t i = 0; i < n; i++) {
      int c = op_pow<CAP>(G, s, n, i, in, out, exp, el, per_elem != 0, expneg, shared_neg != 0);
      if (c && !rc) rc = c;
    }
  })
  return rc;
}

}  // extern "C"

// float codec (cofhe_b200/csrc/codec.cuh) on the host: bit patterns in, sign-magnitude limbs out, and back
extern "C" void hs_codec_make(int k, size_t n, const unsigned *bits, int nl, unsigned *mag /* [n][nl] */, unsigned char *sign) {
  for (size_t i = 0; i < n; i++) {
    float x;
    memcpy(&x, &bits[i], 4);
    sign[i] = (unsigned char)cofhe::codec_make_plaintext(x, k, mag + i * nl, nl);
  }
}
extern "C" void hs_codec_get(int k, size_t n, int nl, const unsigned *mag /* [n][nl] */, const unsigned char *sign, unsigned *bits) {
  for (size_t i = 0; i < n; i++) {
    float f = cofhe::codec_get_float(mag + i * nl, nl, sign[i], k);
    memcpy(&bits[i], &f, 4);
  }
}
