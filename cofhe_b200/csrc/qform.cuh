// cofhe_b200 -- binary quadratic forms of negative discriminant on the generic integer layer:
// canonical reduction, composition with partial reduction (NUCOMP/NUDUPL), exponentiation.
//
// Semantics follow the reference's QFI / ClassGroup (thirdparty/bicycl/src/bicycl/qfi.inl): every public
// operation returns THE reduced form of its class (qfi.inl:523-548), so results are bit-identical to
// ClassGroup::nucomp / nucompinv / nudupl / nupow (qfi.inl:1441-1477) although the computation is organised
// differently (see DESIGN.md "Composition").
#pragma once
#include "bn.cuh"
#include "lehmer.cuh"

namespace cofhe {

template <int CAP>
struct Form {
  Z<CAP> a, b, c;
};

// what the form layer needs to know about the class group
template <int CAP>
struct Group {
  Z<CAP> absD;  // |Delta|
  int dbits;    // bit length of |Delta|
};

// -a < b <= a  (qfi.inl:473-485): b <- b + 2a*q', c <- c + q'*(b + a*q'), q' = floor((a-b)/(2a))
template <int CAP>
QF_HDN void qf_normalize(Err &e, Form<CAP> &f) {
  // already normal when |b| < a, or b == a  (single exit: see DESIGN.md "device-code rules")
  int c = z_cmpabs(f.b, f.a);
  if (!(c < 0 || (c == 0 && !f.b.neg))) {
    Z<CAP> t, q, r, a2;
    z_sub(e, t, f.a, f.b);       // a - b
    z_shl(e, a2, f.a, 1);        // 2a
    z_fdiv_qr<CAP>(e, q, r, t, a2);  // t = q*2a + r, 0 <= r < 2a
    // c += q*(b + a*q)
    z_mul(e, t, f.a, q);
    z_add(e, t, t, f.b);
    z_mul(e, a2, q, t);
    z_add(e, f.c, f.c, a2);
    // b = a - r
    z_sub(e, f.b, f.a, r);
  }
}

template <int CAP>
QF_HDN void qf_reduce(Err &e, Form<CAP> &f) {
  qf_normalize(e, f);
  int guard = 0;
  while (z_cmp(f.a, f.c) > 0 && !e.code) {
    // rho: (a, b, c) <- (c, -b, a), then normalise (qfi.inl:487-492)
    Z<CAP> t;
    z_copy(e, t, f.a);
    z_copy(e, f.a, f.c);
    z_copy(e, f.c, t);
    z_neg(f.b);
    qf_normalize(e, f);
    if (++guard > 64 * CAP) { e.code = 3; break; }
  }
  if (z_cmp(f.a, f.c) == 0 && f.b.neg) f.b.neg = 0;
}

// neutral form (qfi.inl:1333-1345): (1, D mod 2, ((D mod 2) - D)/4)
template <int CAP>
QF_HDN void qf_identity(Err &e, const Group<CAP> &G, Form<CAP> &f) {
  z_set_u32(f.a, 1);
  u32 odd = G.absD.d[0] & 1;
  z_set_u32(f.b, odd);
  Z<CAP> t;
  z_set_u32(t, odd);
  z_add(e, t, t, G.absD);
  z_shr(e, f.c, t, 2);
}

// QFI::neg (qfi.inl:171-176): (a,-b,c) unless that would not be reduced (a == c or a == b)
template <int CAP>
QF_HDN void qf_inverse(Form<CAP> &f) {
  if (!(z_cmp(f.a, f.c) == 0 || z_cmp(f.a, f.b) == 0)) z_neg(f.b);
}

// r = f1 o f2 (neg2: f1 o f2^-1), reduced.  Inputs must be forms of discriminant -absD; they need not be reduced,
// but the partial-reduction bound is tuned for reduced inputs.  dupl: f2 is f1 (skips the first GCD).
// r may alias f1 or f2.
//
// Method (van der Poorten's matrix view, cf. qfi.inl:620-680).  Find a 2x4 integer matrix (A B C D) whose 2x2 minors
// are AB = a1, AC = a2, AD = s, BC = m, BD = c2, CD = c1 with s = (b1+b2)/2, m = (b2-b1)/2; any SL2-multiple of
// it yields an equivalent form
//     ( By*Cy - Ay*Dy ,  (Ax*Dy + Ay*Dx) - (Bx*Cy + By*Cx) ,  Bx*Cx - Ax*Dx ).
// Start from rows x = (G, Bx, ., .), y = (0, a1/G, a2/G, s/G) with F = gcd(a1,a2) = .. + v*a2, G = gcd(F,s) = x0*F + y0*s,
// Bx = (x0*v*m + y0*c2) mod (a1/G); run the Euclidean sequence on the (A,B) columns until B drops below the
// partial-reduction bound; recover the C and D entries of both rows from the orthogonality relations
//     a1*C = a2*B - m*A ,   a1*D = s*B - c2*A .
template <int CAP>
QF_HDN void qf_compose(Err &e, const Group<CAP> &G, Form<CAP> &r, const Form<CAP> &f1, const Form<CAP> &f2, bool neg2,
                       bool dupl) {
  Z<CAP> a1, b1, a2, b2, c2;
  {
    bool sw = z_cmp(f1.a, f2.a) < 0;
    const Form<CAP> &p = sw ? f2 : f1, &q = sw ? f1 : f2;
    z_copy(e, a1, p.a); z_copy(e, b1, p.b);
    z_copy(e, a2, q.a); z_copy(e, b2, q.b); z_copy(e, c2, q.c);
    if (neg2) { if (sw) z_neg(b1); else z_neg(b2); }
  }
  Z<CAP> s, m, t, u;
  z_add(e, t, b1, b2);
  z_shr(e, s, t, 1);
  z_sub(e, m, b2, s);

  Z<CAP> R0, R1, C0, C1;
  Z<CAP> Gg, By, Bx;
  // F = gcd(a1, a2), v*a2 == F (mod a1)
  Z<CAP> F, v;
  if (dupl) {
    z_copy(e, F, a1);
    z_zero(v);
  } else {
    z_copy(e, R0, a1); z_copy(e, R1, a2); z_zero(C0); z_set_u32(C1, 1);
    int st = z_hxgcd(e, R0, R1, C0, C1, 0);
    z_copy(e, F, R0);
    z_copy(e, v, C0);
    if (v.n && (st & 1) == 0) v.neg = 1;  // index of C0 is st; sign(C_i) = (-1)^(i+1)
  }
  if (z_is_one(F)) {
    z_set_u32(Gg, 1);
    z_copy(e, By, a1);
    z_mul(e, t, v, m);
    z_mod<CAP>(e, Bx, t, By);
  } else {
    Z<CAP> sm, x0, y0;
    z_mod<CAP>(e, sm, s, F);
    if (sm.n == 0) {
      z_copy(e, Gg, F);
      z_set_u32(x0, 1);
      z_zero(y0);
    } else {
      z_copy(e, R0, F); z_copy(e, R1, sm); z_zero(C0); z_set_u32(C1, 1);
      int st = z_hxgcd(e, R0, R1, C0, C1, 0);
      z_copy(e, Gg, R0);
      z_copy(e, y0, C0);
      if (y0.n && (st & 1) == 0) y0.neg = 1;
      // x0 = (G - y0*s)/F
      z_mul(e, t, y0, s);
      z_sub(e, t, Gg, t);
      z_divexact<CAP>(e, x0, t, F);
    }
    if (z_is_one(Gg)) z_copy(e, By, a1); else z_divexact<CAP>(e, By, a1, Gg);
    // Bx = (x0*(v*m mod By) + y0*(c2 mod By)) mod By
    z_mul(e, t, v, m);
    z_mod<CAP>(e, u, t, By);
    z_mul(e, t, x0, u);
    z_mod<CAP>(e, Bx, t, By);
    z_mod<CAP>(e, u, c2, By);
    z_mul(e, t, y0, u);
    z_add(e, t, t, Bx);
    z_mod<CAP>(e, Bx, t, By);
  }

  // partial Euclid on (By, Bx): stop near sqrt(a1/a2)*|D|^(1/4)
  int stop = (z_bits(a1) - z_bits(a2)) / 2 + G.dbits / 4;
  if (stop < 1) stop = 1;
  z_copy(e, R0, By); z_copy(e, R1, Bx); z_zero(C0); z_set_u32(C1, 1);
  int st = z_hxgcd(e, R0, R1, C0, C1, stop);
  int idx = 1 + st;
  // rows: x' = (Ax, Bx), y' = (Ay, By)
  Z<CAP> Ax, Ay;
  z_mul(e, Ax, Gg, C0);
  z_mul(e, Ay, Gg, C1);
  if ((idx & 1) == 0) z_neg(Ay);
  z_copy(e, Bx, R0);
  if (idx & 1) z_neg(Bx);
  z_copy(e, By, R1);

  // orthogonality: C = (a2*B - m*A)/a1, D = (s*B - c2*A)/a1
  Z<CAP> Cx, Cy, Dx, Dy;
  z_mul(e, t, a2, By); z_mul(e, u, m, Ay); z_sub(e, t, t, u); z_divexact<CAP>(e, Cy, t, a1);
  z_mul(e, t, s, By);  z_mul(e, u, c2, Ay); z_sub(e, t, t, u); z_divexact<CAP>(e, Dy, t, a1);
  z_mul(e, t, a2, Bx); z_mul(e, u, m, Ax); z_sub(e, t, t, u); z_divexact<CAP>(e, Cx, t, a1);
  z_mul(e, t, s, Bx);  z_mul(e, u, c2, Ax); z_sub(e, t, t, u); z_divexact<CAP>(e, Dx, t, a1);

  z_mul(e, t, By, Cy); z_mul(e, u, Ay, Dy); z_sub(e, r.a, t, u);
  z_mul(e, t, Bx, Cx); z_mul(e, u, Ax, Dx); z_sub(e, r.c, t, u);
  z_mul(e, t, Ax, Dy); z_mul(e, u, Ay, Dx); z_add(e, s, t, u);
  z_mul(e, t, Bx, Cy); z_mul(e, u, By, Cx); z_add(e, t, t, u);
  z_sub(e, r.b, s, t);
  qf_reduce(e, r);
}

// r = f^n for |n| given as nl limbs (little endian) and a sign; n == 0 gives the neutral form, n < 0 applies
// QFI::neg to f^|n| (qfi.inl:977-1064).  Left-to-right binary; only the result is contractual.
template <int CAP>
QF_HDN void qf_pow(Err &e, const Group<CAP> &G, Form<CAP> &r, const Form<CAP> &f, const u32 *n, int nl, size_t stride,
                   bool nneg) {
  int top = -1;
  for (int i = 0; i < nl; i++)
    if (n[(size_t)i * stride] != 0) top = i;
  if (top >= 0) {
    Form<CAP> base;
    z_copy(e, base.a, f.a); z_copy(e, base.b, f.b); z_copy(e, base.c, f.c);
    z_copy(e, r.a, f.a); z_copy(e, r.b, f.b); z_copy(e, r.c, f.c);
    int bit = 31 - clz32(n[(size_t)top * stride]);
    for (int i = 32 * top + bit - 1; i >= 0 && !e.code; i--) {
      qf_compose(e, G, r, r, r, false, true);
      if ((n[(size_t)(i >> 5) * stride] >> (i & 31)) & 1) qf_compose(e, G, r, r, base, false, false);
    }
    if (nneg) qf_inverse(r);
  } else {
    qf_identity(e, G, r);
  }
}

}  // namespace cofhe
