// cofhe_b200 -- cooperative multi-precision integers: one integer spread over the T lanes of a lane group (T = 2, 4,
// 8, ... lanes of one warp), L limbs per lane in registers, K = T*L limbs in all.  Lane j owns limbs [j*L, (j+1)*L).
//
// This is the layer under the cooperative form kernels (coop_qf.cuh) that serve the large discriminants (T1: 2053-bit
// Delta, 33-limb coefficients), where one thread can no longer hold a composition in registers, and the small
// batches, where one thread per form leaves the machine empty and a composition is one warp's latency.
//   * carry chains run inside a lane (IMAD.WIDE / IADD3.X over L limbs); what crosses a lane boundary is ONE signed
//     carry word per lane, moved by a warp shuffle and absorbed by g_propagate (a ripple beyond the neighbour is
//     detected by a group vote and is a 2^-32L event);
//   * products and exact / Euclidean divisions stream one operand limb per step through a sliding window
//     (accumulator words move one limb down per step by a shuffle), operands and results staged in the group's
//     shared-memory scratch so that the stepping loops stay rolled: the code of a composition is a few thousand
//     instructions for every limb count, instead of ~20 k unrolled ones per instantiation of the register path;
//   * scalar control (Lehmer steps, quotient estimates) is computed redundantly by the T lanes of a group.
// Control flow is uniform inside a group; groups of one warp may diverge (shuffles and votes carry the group mask).
//
// The same source compiles for the GPU (nvcc: shuffles, votes, __syncwarp, shared memory) and for the host test
// build (g++: tests/hostsim runs the T lanes of a group as fibers that meet at every exchange).
#pragma once
#include "qf_fast.cuh"

#if defined(__CUDACC__)
#define CG_FN __device__ __forceinline__
#define CG_NOINLINE __device__ __noinline__
#else
#define CG_FN inline
#define CG_NOINLINE
extern "C" void cofhe_sim_gather(void *sim, int lane, unsigned v, unsigned *out);
#endif

namespace cofhe {

template <int T>
struct Grp {
  int lane;   // 0 .. T-1
  u32 *scr;   // scratch of the group (shared memory), cg_scratch_words<T, L>() words
#if defined(__CUDACC__)
  unsigned mask;  // the group's lanes inside the warp
  int base;       // first lane of the group inside the warp
#else
  void *sim;
#endif
};

// scratch map of a group (words): A | B | C | OUT(2K + L)
template <int T, int L>
struct CgScratch { static constexpr int words = 5 * T * L + L; };
template <int T, int L> CG_FN u32 *scr_a(const Grp<T> &g) { return g.scr; }
template <int T, int L> CG_FN u32 *scr_b(const Grp<T> &g) { return g.scr + T * L; }
template <int T, int L> CG_FN u32 *scr_c(const Grp<T> &g) { return g.scr + 2 * T * L; }
template <int T, int L> CG_FN u32 *scr_out(const Grp<T> &g) { return g.scr + 3 * T * L; }

// ---- lane exchange ---------------------------------------------------------------------------------------------
#if defined(__CUDACC__)
template <int T> CG_FN u32 g_bcast(const Grp<T> &g, u32 v, int src) { return T == 1 ? v : __shfl_sync(g.mask, v, src, T); }
template <int T> CG_FN u32 g_up(const Grp<T> &g, u32 v, u32 fill) {
  if (T == 1) return fill;
  u32 t = __shfl_up_sync(g.mask, v, 1, T);
  return g.lane ? t : fill;
}
template <int T> CG_FN u32 g_down(const Grp<T> &g, u32 v, u32 fill) {
  if (T == 1) return fill;
  u32 t = __shfl_down_sync(g.mask, v, 1, T);
  return g.lane == T - 1 ? fill : t;
}
template <int T> CG_FN u32 g_ballot(const Grp<T> &g, bool p) {
  if (T == 1) return p ? 1u : 0u;
  return (__ballot_sync(g.mask, p) & g.mask) >> g.base;
}
template <int T> CG_FN void g_sync(const Grp<T> &g) { if (T > 1) __syncwarp(g.mask); }
#else
template <int T> CG_FN u32 g_bcast(const Grp<T> &g, u32 v, int src) { u32 o[32]; cofhe_sim_gather(g.sim, g.lane, v, o); return o[src]; }
template <int T> CG_FN u32 g_up(const Grp<T> &g, u32 v, u32 fill) { u32 o[32]; cofhe_sim_gather(g.sim, g.lane, v, o); return g.lane ? o[g.lane - 1] : fill; }
template <int T> CG_FN u32 g_down(const Grp<T> &g, u32 v, u32 fill) { u32 o[32]; cofhe_sim_gather(g.sim, g.lane, v, o); return g.lane == T - 1 ? fill : o[g.lane + 1]; }
template <int T> CG_FN u32 g_ballot(const Grp<T> &g, bool p) {
  u32 o[32], r = 0;
  cofhe_sim_gather(g.sim, g.lane, p ? 1u : 0u, o);
  for (int i = 0; i < T; i++) r |= (o[i] & 1u) << i;
  return r;
}
template <int T> CG_FN void g_sync(const Grp<T> &g) { u32 o[32]; cofhe_sim_gather(g.sim, g.lane, 0, o); }
#endif
template <int T> CG_FN bool g_any(const Grp<T> &g, bool p) { return g_ballot<T>(g, p) != 0; }

// ---- carries across lanes ----------------------------------------------------------------------------------------
// Every lane has produced its block x (L limbs, modulo 2^32L) and a signed word `cw` that belongs to the block of
// the lane above.  Adds the words in, rippling until nothing is left.  Returns, in lane T-1 (0 elsewhere), what
// left the top block.  WIDE: the first words may need more than 32 bits (products by a full limb).
template <int T, int L, bool WIDE>
CG_FN i64 g_propagate(const Grp<T> &g, u32 *x, i64 cw) {
  i64 top = 0;
  for (;;) {
    if (g.lane == T - 1) top += cw;
    i64 c;
    if (WIDE) {
      u32 lo = g_up<T>(g, (u32)cw, 0), hi = g_up<T>(g, (u32)((u64)cw >> 32), 0);
      c = (i64)(((u64)hi << 32) | lo);
    } else {
      c = (i64)(int)g_up<T>(g, (u32)cw, 0);
    }
    QF_UNROLL for (int i = 0; i < L; i++) { c += x[i]; x[i] = (u32)c; c >>= 32; }
    cw = c;
    if (!g_any<T>(g, cw != 0)) break;
  }
  return top;
}
// value of lane T-1's i64 in every lane
template <int T> CG_FN i64 g_top64(const Grp<T> &g, i64 v) {
  u32 lo = g_bcast<T>(g, (u32)v, T - 1), hi = g_bcast<T>(g, (u32)((u64)v >> 32), T - 1);
  return (i64)(((u64)hi << 32) | lo);
}

// r = a + b; returns the carry out (every lane)
template <int T, int L>
CG_FN u32 g_add(const Grp<T> &g, u32 *r, const u32 *a, const u32 *b) {
  u32 c = fw_add<L>(r, a, b);
  i64 top = g_propagate<T, L, false>(g, r, (i64)c);
  return g_bcast<T>(g, (u32)top, T - 1);
}
// r = a - b; returns the borrow out (1 if a < b as unsigned), every lane
template <int T, int L>
CG_FN u32 g_sub(const Grp<T> &g, u32 *r, const u32 *a, const u32 *b) {
  u32 bo = fw_sub<L>(r, a, b);
  i64 top = g_propagate<T, L, false>(g, r, -(i64)bo);
  return g_bcast<T>(g, (u32)(-top), T - 1);
}
// same, results not needed (two's complement arithmetic)
template <int T, int L> CG_FN void g_addm(const Grp<T> &g, u32 *r, const u32 *a, const u32 *b) {
  u32 c = fw_add<L>(r, a, b);
  g_propagate<T, L, false>(g, r, (i64)c);
}
template <int T, int L> CG_FN void g_subm(const Grp<T> &g, u32 *r, const u32 *a, const u32 *b) {
  u32 bo = fw_sub<L>(r, a, b);
  g_propagate<T, L, false>(g, r, -(i64)bo);
}
template <int T, int L>
CG_FN void g_neg(const Grp<T> &g, u32 *r, const u32 *a) {
  u64 c = g.lane == 0 ? 1 : 0;
  QF_UNROLL for (int i = 0; i < L; i++) { c += (u32)~a[i]; r[i] = (u32)c; c >>= 32; }
  g_propagate<T, L, false>(g, r, (i64)c);
}
template <int T, int L>
CG_FN void g_cneg(const Grp<T> &g, u32 *r, const u32 *a, bool neg) {  // neg is uniform in the group
  if (neg) g_neg<T, L>(g, r, a);
  else { QF_UNROLL for (int i = 0; i < L; i++) r[i] = a[i]; }
}
template <int T, int L> CG_FN void g_copy(u32 *r, const u32 *a) { QF_UNROLL for (int i = 0; i < L; i++) r[i] = a[i]; }
template <int T, int L> CG_FN void g_zero(u32 *r) { QF_UNROLL for (int i = 0; i < L; i++) r[i] = 0; }
template <int T, int L> CG_FN void g_set_u32(const Grp<T> &g, u32 *r, u32 v) {
  QF_UNROLL for (int i = 0; i < L; i++) r[i] = 0;
  if (g.lane == 0) r[0] = v;
}

// ---- predicates and sizes (results uniform in the group) ---------------------------------------------------------
template <int T, int L> CG_FN bool g_is_zero(const Grp<T> &g, const u32 *a) { return !g_any<T>(g, !fw_is_zero<L>(a)); }
template <int T, int L>
CG_FN int g_cmp(const Grp<T> &g, const u32 *a, const u32 *b) {  // unsigned: -1, 0, 1
  int lc = fw_cmp<L>(a, b);
  u32 ne = g_ballot<T>(g, lc != 0);
  if (!ne) return 0;
  return (int)g_bcast<T>(g, (u32)lc, 31 - clz32(ne));
}
template <int T, int L>
CG_FN int g_bits(const Grp<T> &g, const u32 *a) {
  int lb = fw_bits<L>(a);
  u32 nz = g_ballot<T>(g, lb != 0);
  if (!nz) return 0;
  int tl = 31 - clz32(nz);
  return 32 * L * tl + (int)g_bcast<T>(g, (u32)lb, tl);
}
template <int T, int L>
CG_FN int g_ctz(const Grp<T> &g, const u32 *a) {  // a != 0
  bool z = fw_is_zero<L>(a);
  int lc = z ? 0 : fw_ctz<L>(a);
  u32 nz = g_ballot<T>(g, !z);
  if (!nz) return 32 * T * L;
  int ll = ctz32(nz);
  return 32 * L * ll + (int)g_bcast<T>(g, (u32)lc, ll);
}
// sign bit of a two's complement value
template <int T, int L> CG_FN bool g_sign(const Grp<T> &g, const u32 *a) { return (g_bcast<T>(g, a[L - 1], T - 1) >> 31) != 0; }
// limb `idx` (dynamic, uniform) of a
template <int T, int L>
CG_FN u32 g_limb(const Grp<T> &g, const u32 *a, int idx) {
  int li = idx % L;
  u32 v = 0;
  QF_UNROLL for (int i = 0; i < L; i++) v = (i == li) ? a[i] : v;
  return g_bcast<T>(g, v, idx / L);
}
// limbs >= m (static) replaced by `fill`
template <int T, int L, int M_>
CG_FN void g_fill_from(const Grp<T> &g, u32 *a, u32 fill) {
  QF_UNROLL for (int i = 0; i < L; i++) a[i] = (g.lane * L + i >= M_) ? fill : a[i];
}
// are limbs >= m (dynamic) all zero?
template <int T, int L>
CG_FN bool g_high_zero(const Grp<T> &g, const u32 *a, int m) {
  u32 o = 0;
  QF_UNROLL for (int i = 0; i < L; i++) o |= (g.lane * L + i >= m) ? a[i] : 0u;
  return !g_any<T>(g, o != 0);
}

// ---- shifts -------------------------------------------------------------------------------------------------------
template <int T, int L>
CG_FN void g_shl_bits(const Grp<T> &g, u32 *r, const u32 *a, int s) {  // 0 <= s < 32
  u32 below = g_up<T>(g, a[L - 1], 0);
  if (s == 0) { g_copy<T, L>(r, a); return; }
  QF_UNROLL for (int i = L - 1; i > 0; i--) r[i] = (a[i] << s) | (a[i - 1] >> (32 - s));
  r[0] = (a[0] << s) | (below >> (32 - s));
}
template <int T, int L>
CG_FN void g_shr_bits(const Grp<T> &g, u32 *r, const u32 *a, int s, u32 fill = 0) {  // fill: the limb above the top
  u32 above = g_down<T>(g, a[0], fill);
  if (s == 0) { g_copy<T, L>(r, a); return; }
  QF_UNROLL for (int i = 0; i < L - 1; i++) r[i] = (a[i] >> s) | (a[i + 1] << (32 - s));
  r[L - 1] = (a[L - 1] >> s) | (above << (32 - s));
}
// whole-limb shifts by a dynamic (uniform) count through the scratch slot C
template <int T, int L>
CG_FN void g_shl_limbs(const Grp<T> &g, u32 *x, int k) {
  if (k == 0) return;
  u32 *s = scr_c<T, L>(g);
  g_sync<T>(g);
  QF_UNROLL for (int i = 0; i < L; i++) s[g.lane * L + i] = x[i];
  g_sync<T>(g);
  QF_UNROLL for (int i = 0; i < L; i++) { int p = g.lane * L + i - k; x[i] = p >= 0 ? s[p] : 0u; }
}
template <int T, int L>
CG_FN void g_shr_limbs(const Grp<T> &g, u32 *x, int k, u32 fill = 0) {
  if (k == 0) return;
  u32 *s = scr_c<T, L>(g);
  g_sync<T>(g);
  QF_UNROLL for (int i = 0; i < L; i++) s[g.lane * L + i] = x[i];
  g_sync<T>(g);
  QF_UNROLL for (int i = 0; i < L; i++) { int p = g.lane * L + i + k; x[i] = p < T * L ? s[p] : fill; }
}
// x <- x * 2^32 + w  (one limb up, w uniform); returns nothing, the limb that falls off the top is lost
template <int T, int L>
CG_FN void g_slide_up(const Grp<T> &g, u32 *x, u32 w) {
  u32 in = g_up<T>(g, x[L - 1], w);
  QF_UNROLL for (int i = L - 1; i > 0; i--) x[i] = x[i - 1];
  x[0] = in;
}
// block of scratch -> registers and back (lane-blocked)
template <int T, int L> CG_FN void g_from_scr(const Grp<T> &g, u32 *x, const u32 *s) { QF_UNROLL for (int i = 0; i < L; i++) x[i] = s[g.lane * L + i]; }
template <int T, int L> CG_FN void g_to_scr(const Grp<T> &g, u32 *s, const u32 *x) { QF_UNROLL for (int i = 0; i < L; i++) s[g.lane * L + i] = x[i]; }

// ---- products by one limb and linear combinations ------------------------------------------------------------------
// r = a * w; returns the limb above the top (every lane)
template <int T, int L>
CG_FN u32 g_mul1(const Grp<T> &g, u32 *r, const u32 *a, u32 w) {
  u64 c = 0;
  QF_UNROLL for (int i = 0; i < L; i++) { c += (u64)a[i] * w; r[i] = (u32)c; c >>= 32; }
  i64 top = g_propagate<T, L, true>(g, r, (i64)c);
  return g_bcast<T>(g, (u32)top, T - 1);
}
// x -= a * w; returns what had to be borrowed from above the top (every lane; up to 33 bits)
template <int T, int L>
CG_FN u64 g_submul1(const Grp<T> &g, u32 *x, const u32 *a, u32 w) {
  u64 c = 0;
  QF_UNROLL for (int i = 0; i < L; i++) {
    u64 p = (u64)a[i] * w + c;
    u32 lo = (u32)p;
    c = (p >> 32) + (x[i] < lo ? 1 : 0);
    x[i] -= lo;
  }
  i64 top = g_propagate<T, L, true>(g, x, -(i64)c);
  return (u64)(-g_top64<T>(g, top));
}
// r = |x*A - y*B| with the sign known (neg: the combination is <= 0); x, y < 2^30
template <int T, int L>
CG_FN void g_lin_sub(const Grp<T> &g, u32 *r, u32 x, const u32 *A, u32 y, const u32 *B, bool neg) {
  i64 c = 0;
  u32 p = neg ? y : x, q = neg ? x : y;
  const u32 *P = neg ? B : A, *Q = neg ? A : B;
  QF_UNROLL for (int i = 0; i < L; i++) {
    c += (i64)((u64)p * P[i]) - (i64)((u64)q * Q[i]);
    r[i] = (u32)c;
    c >>= 32;
  }
  g_propagate<T, L, false>(g, r, c);
}
// r = x*A + y*B; returns the limb above the top (every lane).  x, y < 2^30
template <int T, int L>
CG_FN u32 g_lin_add(const Grp<T> &g, u32 *r, u32 x, const u32 *A, u32 y, const u32 *B) {
  u64 c = 0;
  QF_UNROLL for (int i = 0; i < L; i++) {
    c += (u64)x * A[i] + (u64)y * B[i];
    r[i] = (u32)c;
    c >>= 32;
  }
  i64 top = g_propagate<T, L, false>(g, r, (i64)c);
  return g_bcast<T>(g, (u32)top, T - 1);
}

// ---- division by one limb (lane-serial chains: the remainder / borrow travels from lane to lane) ------------------
// a mod w, w != 0
template <int T, int L>
CG_FN u32 g_mod1(const Grp<T> &g, const u32 *a, u32 w) {
  int s = clz32(w);
  u32 d = w << s, v = recip32(d);
  u32 below = g_up<T>(g, a[L - 1], 0);
  u32 r = s ? (g_bcast<T>(g, a[L - 1], T - 1) >> (32 - s)) : 0u;
  for (int j = T - 1; j >= 0; j--) {
    u32 rr = r;
    QF_UNROLL for (int i = L - 1; i >= 0; i--) {
      u32 lo = s ? ((a[i] << s) | ((i ? a[i > 0 ? i - 1 : 0] : below) >> (32 - s))) : a[i];
      u32 rem;
      div_2by1(rr, lo, d, v, &rem);
      rr = rem;
    }
    r = g_bcast<T>(g, rr, j);
  }
  return r >> s;
}
// q = a / w exactly (w != 0, w | a)
template <int T, int L>
CG_FN void g_divexact1(const Grp<T> &g, u32 *q, const u32 *a, u32 w) {
  int t = ctz32(w);
  u32 wo = w >> t, wi = inv32(wo);
  u32 x[L];
  g_shr_bits<T, L>(g, x, a, t);
  u32 bor = 0;
  for (int j = 0; j < T; j++) {
    u32 b = bor, ql[L];
    QF_UNROLL for (int i = 0; i < L; i++) {
      u32 cur = x[i] - b;
      u32 b1 = x[i] < b ? 1u : 0u;
      u32 qi = cur * wi;
      ql[i] = qi;
      b = (u32)(((u64)qi * wo) >> 32) + b1;
    }
    if (g.lane == j) { QF_UNROLL for (int i = 0; i < L; i++) q[i] = ql[i]; }
    bor = g_bcast<T>(g, b, j);
  }
}

// ---- streaming operations on the scratch ---------------------------------------------------------------------------
// out[0 .. K + nbr) = A * B[0 .. nb)   (unsigned; nbr = nb rounded up to a multiple of L; A: K limbs, B: >= nbr readable
// limbs of which those >= nb are ignored; out must not overlap A or B).  Sliding window: step k adds A * B[k] into an
// accumulator whose word 0 is then final (it is product limb k, stored by lane 0) and moves one limb down.
template <int T, int L>
CG_NOINLINE void g_mul_s(const Grp<T> g, const u32 *A, const u32 *B, int nb, u32 *out) {
  u32 a[L], acc[L];
  g_sync<T>(g);
  QF_UNROLL for (int i = 0; i < L; i++) { a[i] = A[g.lane * L + i]; acc[i] = 0; }
  u32 cw = 0;  // pending word for the block above (<= 2)
  int k0 = 0;
  QF_NOUNROLL for (; k0 < nb; k0 += L) {
    QF_UNROLL for (int r = 0; r < L; r++) {
      // window word w lives in acc[(w + r) % L]
      u32 bk = (k0 + r < nb) ? B[k0 + r] : 0u;
      u64 c = 0;
      QF_UNROLL for (int i = 0; i < L; i++) {
        c += (u64)a[i] * bk + acc[(i + r) % L];
        acc[(i + r) % L] = (u32)c;
        c >>= 32;
      }
      if (g.lane == 0) out[k0 + r] = acc[r];
      u32 dn = g_down<T>(g, acc[r], 0);
      u64 t = (u64)dn + cw + c;
      acc[r] = (u32)t;
      cw = (u32)(t >> 32);
    }
  }
  g_propagate<T, L, false>(g, acc, (i64)cw);
  QF_UNROLL for (int i = 0; i < L; i++) out[k0 + g.lane * L + i] = acc[i];
  g_sync<T>(g);
}

// Q[0 .. mr) = NUM / D in Z / 2^(32 mr) for odd D (Jebelean's exact division, least significant limb first);
// mr = m rounded up to a multiple of L.  NUM, D: K limbs.  Q must not overlap NUM or D.
template <int T, int L>
CG_NOINLINE void g_divexact2_s(const Grp<T> g, const u32 *NUM, const u32 *D, int m, u32 *Q) {
  u32 num[L], d[L];
  g_sync<T>(g);
  QF_UNROLL for (int i = 0; i < L; i++) { num[i] = NUM[g.lane * L + i]; d[i] = D[g.lane * L + i]; }
  const u32 dinv = inv32(D[0]);
  u32 bw = 0;  // pending borrow for the block above (<= 2)
  QF_NOUNROLL for (int k0 = 0; k0 < m; k0 += L) {
    QF_UNROLL for (int r = 0; r < L; r++) {
      u32 q = g_bcast<T>(g, num[r], 0) * dinv;
      if (g.lane == 0) Q[k0 + r] = q;
      u64 c = 0;
      QF_UNROLL for (int i = 0; i < L; i++) {
        u64 p = (u64)q * d[i] + c;
        u32 lo = (u32)p;
        c = (p >> 32) + (num[(i + r) % L] < lo ? 1 : 0);
        num[(i + r) % L] -= lo;
      }
      u32 dn = g_down<T>(g, num[r], 0);
      i64 t = (i64)(u64)dn - (i64)c - (i64)bw;
      num[r] = (u32)t;
      bw = (u32)(-(t >> 32));
    }
  }
  g_sync<T>(g);
}

// limb p of (X << (32*ls + s)) for X[0 .. nx) in memory
CG_FN u32 sh_limb(const u32 *X, int nx, int p, int ls, int s) {
  int q = p - ls;
  u32 hi = (q >= 0 && q < nx) ? X[q] : 0u, lo = (q - 1 >= 0 && q - 1 < nx) ? X[q - 1] : 0u;
  return s ? ((hi << s) | (lo >> (32 - s))) : hi;
}

// R = U mod D.  U: nu limbs, D: K limbs (both in scratch), R: K limbs of scratch (may not overlap U or D; slot C is
// used inside).  Knuth D on a left-justified divisor; the dividend enters one limb per step at the bottom of a
// K-limb window.  Returns false when D == 0.
template <int T, int L>
CG_NOINLINE bool g_mod_s(const Grp<T> g, const u32 *U, int nu, const u32 *D, u32 *R) {
  constexpr int K = T * L;
  g_sync<T>(g);
  int nd = K;
  while (nd > 0 && D[nd - 1] == 0) nd--;
  if (nd == 0) return false;
  while (nu > 0 && U[nu - 1] == 0) nu--;
  const int s = clz32(D[nd - 1]), ls = K - nd;
  // U' = U << (32 ls + s) has nup limbs; D' = D << (32 ls + s) fills K limbs with its top bit set
  const int nup = nu == 0 ? 0 : nu + ls + ((s && (U[nu - 1] >> (32 - s))) ? 1 : 0);
  u32 x[L], d[L];
  if (nup < K) {
    // U < D
    QF_UNROLL for (int i = 0; i < L; i++) { int p = g.lane * L + i; x[i] = p < nu ? U[p] : 0u; }
    g_sync<T>(g);
    g_to_scr<T, L>(g, R, x);
    g_sync<T>(g);
    return true;
  }
  QF_UNROLL for (int i = 0; i < L; i++) {
    d[i] = sh_limb(D, nd, g.lane * L + i, ls, s);
    x[i] = sh_limb(U, nu, nup - K + g.lane * L + i, ls, s);
  }
  if (g_cmp<T, L>(g, x, d) >= 0) g_subm<T, L>(g, x, x, d);
  const u32 dt = sh_limb(D, nd, K - 1, ls, s), v = recip32(dt);
  QF_NOUNROLL for (int p = nup - K - 1; p >= 0; p--) {
    u32 nxt = sh_limb(U, nu, p, ls, s);
    u32 rt = g_bcast<T>(g, x[L - 1], T - 1);
    u32 t0 = L >= 2 ? g_bcast<T>(g, x[L >= 2 ? L - 2 : 0], T - 1) : g_bcast<T>(g, x[0], T >= 2 ? T - 2 : 0);
    g_slide_up<T, L>(g, x, nxt);
    // (rt : x) / d, with rt <= dt
    u32 rem;
    u32 qh = (rt >= dt) ? 0xffffffffu : div_2by1(rt, t0, dt, v, &rem);
    if (qh != 0) {
      u64 bo = g_submul1<T, L>(g, x, d, qh);
      if ((u64)rt < bo) {
        // went negative: add back (at most twice)
        u32 top = rt - (u32)bo;
        top += g_add<T, L>(g, x, x, d);
        if (top != 0) g_add<T, L>(g, x, x, d);
      }
    }
  }
  // remainder = x >> (32 ls + s)
  u32 *tmp = scr_c<T, L>(g);
  g_sync<T>(g);
  g_to_scr<T, L>(g, tmp, x);
  g_sync<T>(g);
  QF_UNROLL for (int i = 0; i < L; i++) {
    int q = g.lane * L + i + ls;
    u32 lo = q < K ? tmp[q] : 0u, hi = q + 1 < K ? tmp[q + 1] : 0u;
    x[i] = s ? ((lo >> s) | (hi << (32 - s))) : lo;
  }
  g_to_scr<T, L>(g, R, x);
  g_sync<T>(g);
  return true;
}

}  // namespace cofhe
