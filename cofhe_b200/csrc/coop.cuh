// cofhe_b200 -- cooperative multi-precision integers: one integer spread over the T lanes of a lane group (T = 2, 4,
// 8, ... lanes of one warp), L limbs per lane in registers, K = T*L limbs in all.  Lane j owns limbs [j*L, (j+1)*L).
//
// This is the layer under the cooperative form kernels (coop_qf.cuh) that serve the large discriminants (T1: 2053-bit
// Delta, 33-limb coefficients), where one thread can no longer hold a composition in registers.
//   * carry chains run inside a lane (IMAD.WIDE / IADD3.X over L limbs); what crosses a lane boundary is ONE signed
//     carry word per lane, moved by a warp shuffle and absorbed by g_propagate (a ripple beyond the neighbour is
//     detected by a vote and is a 2^-32L event);
//   * products and exact / Euclidean divisions stream one operand limb per step through a sliding window
//     (accumulator words move one limb down per step by a shuffle), operands and results staged in the group's
//     shared-memory scratch so that the stepping loops stay rolled: the code of a composition is a few thousand
//     instructions for every limb count, instead of ~20 k unrolled ones per instantiation of the register path;
//   * scalar control (Lehmer steps, quotient estimates) is computed redundantly by the T lanes of a group.
//
// WARP-CONVERGED.  The 32/T groups of a warp run in lock step: every exchange (shuffle, vote, scratch barrier) is
// executed by all 32 lanes with the full mask, so it is ONE instruction for all groups of the warp.  (The first version
// let groups diverge and passed the group's own mask: ptxas then wraps every shuffle in a MATCH / REDUX / VOTE /
// BRA.DIV sequence and the hardware runs the groups one after the other -- 8.6 of 32 lanes active on the shuffles,
// profiles/r2c_ncu_k_coop_compose_t1.txt.)  The price is that control flow must be warp-uniform wherever an exchange
// follows: data-dependent branches are taken when ANY group of the warp needs them (w_any) and the groups that do not
// are carried through with a predicate (`act` arguments, *_if operations), loops run to the warp's maximum trip count.
// A group that has declined its element (bail) keeps executing on whatever values it has; nothing here can trap, loop
// forever or index out of range on arbitrary values.
//
// The same source compiles for the GPU (nvcc: shuffles, votes, __syncwarp, shared memory) and for the host test
// build (g++: tests/hostsim runs the lanes of a warp as fibers that meet at every exchange).
#pragma once
#include "qf_fast.cuh"

#if defined(__CUDACC__)
#define CG_FN __device__ __forceinline__
#if defined(CG_INLINE_ALL)
#define CG_NOINLINE __device__ __forceinline__
#else
#define CG_NOINLINE __device__ __noinline__
#endif
#else
#define CG_FN inline
#define CG_NOINLINE
extern "C" void cofhe_sim_gather(void *sim, int wlane, unsigned v, unsigned *out);
extern "C" int cofhe_sim_width(void *sim);
#endif

// debugging aid (tools/coop_one.cu): progress marks into mapped host memory, readable after a device fault
#if defined(CG_DEBUG) && defined(__CUDACC__)
static __device__ volatile int *cg_dbg = nullptr;
#define CG_MARK(k) do { if (cg_dbg && (threadIdx.x & 31) == 0 && blockIdx.x == 0) { cg_dbg[threadIdx.x >> 5] = (k); __threadfence_system(); } } while (0)
#define CG_NOTE(slot, v) do { if (cg_dbg && threadIdx.x < 32 && g.lane == 0 && blockIdx.x == 0) { cg_dbg[(slot) + g.base / T] = (int)(v); __threadfence_system(); } } while (0)
#else
#define CG_MARK(k) ((void)0)
#define CG_NOTE(slot, v) ((void)0)
#endif

namespace cofhe {

template <int T>
struct Grp {
  int lane;   // 0 .. T-1
  int base;   // first lane of the group inside the warp
  u32 *scr;   // scratch of the group (shared memory), CgScratch<T, L>::words words
#if defined(__CUDACC__)
  unsigned all;  // 0xffffffff, read with __activemask() at kernel entry: the member mask of every vote (see w_any)
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

// ---- lane exchange (all 32 lanes of the warp take part in every one of these) -----------------------------------
#if defined(__CUDACC__)
#define CG_FULL 0xffffffffu
#if defined(CG_DEBUG)
// debug: record (once) the progress mark at which an exchange was first reached by a partial warp
CG_FN void cg_check_converged(int kind) {
  unsigned am = __activemask();
  if (am != CG_FULL && cg_dbg && threadIdx.x < 32 && blockIdx.x == 0 && cg_dbg[48] == -1) {
    cg_dbg[48] = cg_dbg[0]; cg_dbg[49] = (int)am; cg_dbg[50] = kind; __threadfence_system();
  }
}
#else
CG_FN void cg_check_converged(int) {}
#endif
template <int T> CG_FN u32 g_bcast(const Grp<T> &g, u32 v, int src) { cg_check_converged(3); return T == 1 ? v : __shfl_sync(CG_FULL, v, src, T); }
template <int T> CG_FN u32 g_up(const Grp<T> &g, u32 v, u32 fill) {
  if (T == 1) return fill;
  u32 t = __shfl_up_sync(CG_FULL, v, 1, T);
  return g.lane ? t : fill;
}
template <int T> CG_FN u32 g_down(const Grp<T> &g, u32 v, u32 fill) {
  if (T == 1) return fill;
  u32 t = __shfl_down_sync(CG_FULL, v, 1, T);
  return g.lane == T - 1 ? fill : t;
}
template <int T> CG_FN u32 g_ballot(const Grp<T> &g, bool p) {
  cg_check_converged(2);
  u32 b = __ballot_sync(g.all, p);
  return T == 32 ? b : ((b >> g.base) & ((1u << T) - 1u));
}
// Votes and reductions take the member mask from a register (g.all = 0xffffffff, opaque to the compiler).  With the
// literal mask ptxas emits a bare VOTE wherever it considers the warp converged -- e.g. straight after the BSYNC that
// closes the lane-divergent scalar Euclid -- and on sm_100a the lanes of one group were observed to be missing at
// exactly such a vote (tools/coop_one.cu: active mask 0xfff0ffff), after which the groups of the warp fell out of step.
// A register mask makes ptxas guard the vote (MATCH / BRA.DIV, WARPSYNC on the slow path), as it does for every SHFL.
template <int T, int ID = 1> CG_FN bool w_any(const Grp<T> &g, bool p) { cg_check_converged(ID); return __any_sync(g.all, p) != 0; }
template <int T> CG_FN u32 w_max(const Grp<T> &g, u32 v) { return __reduce_max_sync(g.all, v); }
template <int T> CG_FN void g_sync(const Grp<T> &g) { __syncwarp(); }
#else
template <int T> CG_FN u32 g_bcast(const Grp<T> &g, u32 v, int src) { u32 o[32]; cofhe_sim_gather(g.sim, g.base + g.lane, v, o); return o[g.base + src]; }
template <int T> CG_FN u32 g_up(const Grp<T> &g, u32 v, u32 fill) { u32 o[32]; cofhe_sim_gather(g.sim, g.base + g.lane, v, o); return g.lane ? o[g.base + g.lane - 1] : fill; }
template <int T> CG_FN u32 g_down(const Grp<T> &g, u32 v, u32 fill) { u32 o[32]; cofhe_sim_gather(g.sim, g.base + g.lane, v, o); return g.lane == T - 1 ? fill : o[g.base + g.lane + 1]; }
template <int T> CG_FN u32 g_ballot(const Grp<T> &g, bool p) {
  u32 o[32], r = 0;
  cofhe_sim_gather(g.sim, g.base + g.lane, p ? 1u : 0u, o);
  for (int i = 0; i < T; i++) r |= (o[g.base + i] & 1u) << i;
  return r;
}
template <int T, int ID = 1> CG_FN bool w_any(const Grp<T> &g, bool p) {
  u32 o[32], r = 0;
  cofhe_sim_gather(g.sim, g.base + g.lane, p ? 1u : 0u, o);
  for (int i = 0; i < cofhe_sim_width(g.sim); i++) r |= o[i];
  return r != 0;
}
template <int T> CG_FN u32 w_max(const Grp<T> &g, u32 v) {
  u32 o[32], r = 0;
  cofhe_sim_gather(g.sim, g.base + g.lane, v, o);
  for (int i = 0; i < cofhe_sim_width(g.sim); i++) r = o[i] > r ? o[i] : r;
  return r;
}
template <int T> CG_FN void g_sync(const Grp<T> &g) { u32 o[32]; cofhe_sim_gather(g.sim, g.base + g.lane, 0, o); }
#endif
template <int T> CG_FN bool g_any(const Grp<T> &g, bool p) { return g_ballot<T>(g, p) != 0; }

// ---- carries across lanes ----------------------------------------------------------------------------------------
// Every lane has produced its block x (L limbs, modulo 2^32L) and a signed word `cw` that belongs to the block of
// the lane above.  Adds the words in, rippling until nothing is left anywhere in the warp.  Returns, in lane T-1 (0
// elsewhere), what left the top block.  WIDE: the first words may need more than 32 bits (products by a full limb).
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
    // the word lands in limb 0; what it pushes on is -1, 0 or +1 and dies in limb 1 unless that limb is all zeros / all ones
    // (2^-32): the remaining limbs are touched only when some lane of the warp still carries something
    static_assert(L >= 2, "two limbs absorb the word");
    c += x[0]; x[0] = (u32)c; c >>= 32;
    c += x[1]; x[1] = (u32)c; c >>= 32;
    if (!w_any<T, 90>(g, c != 0)) break;
    QF_UNROLL for (int i = 2; i < L; i++) { c += x[i]; x[i] = (u32)c; c >>= 32; }
    cw = c;
    if (!w_any<T, 91>(g, cw != 0)) break;
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
// r = a + (on ? b : 0), r = a - (on ? b : 0): the predicated forms (on is uniform in the group)
template <int T, int L> CG_FN u32 g_add_if(const Grp<T> &g, u32 *r, const u32 *a, const u32 *b, bool on) {
  u32 m = on ? 0xffffffffu : 0u, t[L];
  QF_UNROLL for (int i = 0; i < L; i++) t[i] = b[i] & m;
  return g_add<T, L>(g, r, a, t);
}
template <int T, int L> CG_FN void g_sub_if(const Grp<T> &g, u32 *r, const u32 *a, const u32 *b, bool on) {
  u32 m = on ? 0xffffffffu : 0u, t[L];
  QF_UNROLL for (int i = 0; i < L; i++) t[i] = b[i] & m;
  g_subm<T, L>(g, r, a, t);
}
// r = neg ? -a : a  (neg is uniform in the group; the chain always runs)
template <int T, int L>
CG_FN void g_cneg(const Grp<T> &g, u32 *r, const u32 *a, bool neg) {
  u32 m = neg ? 0xffffffffu : 0u;
  u64 c = (neg && g.lane == 0) ? 1 : 0;
  QF_UNROLL for (int i = 0; i < L; i++) { c += (u32)(a[i] ^ m); r[i] = (u32)c; c >>= 32; }
  g_propagate<T, L, false>(g, r, (i64)c);
}
template <int T, int L> CG_FN void g_neg(const Grp<T> &g, u32 *r, const u32 *a) { g_cneg<T, L>(g, r, a, true); }
template <int T, int L> CG_FN void g_copy(u32 *r, const u32 *a) { QF_UNROLL for (int i = 0; i < L; i++) r[i] = a[i]; }
template <int T, int L> CG_FN void g_copy_if(u32 *r, const u32 *a, bool on) { QF_UNROLL for (int i = 0; i < L; i++) r[i] = on ? a[i] : r[i]; }
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
  int top = ne ? 31 - clz32(ne) : 0;
  int v = (int)g_bcast<T>(g, (u32)lc, top);
  return ne ? v : 0;
}
template <int T, int L>
CG_FN int g_bits(const Grp<T> &g, const u32 *a) {
  int lb = fw_bits<L>(a);
  u32 nz = g_ballot<T>(g, lb != 0);
  int tl = nz ? 31 - clz32(nz) : 0;
  int v = (int)g_bcast<T>(g, (u32)lb, tl);
  return nz ? 32 * L * tl + v : 0;
}
template <int T, int L>
CG_FN int g_ctz(const Grp<T> &g, const u32 *a) {  // 32*K for a == 0
  bool z = fw_is_zero<L>(a);
  int lc = z ? 0 : fw_ctz<L>(a);
  u32 nz = g_ballot<T>(g, !z);
  int ll = nz ? ctz32(nz) : 0;
  int v = (int)g_bcast<T>(g, (u32)lc, ll);
  return nz ? 32 * L * ll + v : 32 * T * L;
}
// sign bit of a two's complement value
template <int T, int L> CG_FN bool g_sign(const Grp<T> &g, const u32 *a) { return (g_bcast<T>(g, a[L - 1], T - 1) >> 31) != 0; }
// limb `idx` (dynamic, uniform in the group, clamped to the width) of a
template <int T, int L>
CG_FN u32 g_limb(const Grp<T> &g, const u32 *a, int idx) {
  idx = idx < 0 ? 0 : (idx >= T * L ? T * L - 1 : idx);
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
// whole-limb shifts by a dynamic count (uniform in the group, 0 <= k) through the scratch slot C
template <int T, int L>
CG_FN void g_shl_limbs(const Grp<T> &g, u32 *x, int k) {
  u32 *s = scr_c<T, L>(g);
  g_sync<T>(g);
  QF_UNROLL for (int i = 0; i < L; i++) s[g.lane * L + i] = x[i];
  g_sync<T>(g);
  QF_UNROLL for (int i = 0; i < L; i++) { int p = g.lane * L + i - k; x[i] = p >= 0 ? s[p] : 0u; }
}
template <int T, int L>
CG_FN void g_shr_limbs(const Grp<T> &g, u32 *x, int k, u32 fill = 0) {
  u32 *s = scr_c<T, L>(g);
  g_sync<T>(g);
  QF_UNROLL for (int i = 0; i < L; i++) s[g.lane * L + i] = x[i];
  g_sync<T>(g);
  QF_UNROLL for (int i = 0; i < L; i++) { int p = g.lane * L + i + k; x[i] = p < T * L ? s[p] : fill; }
}
// x <- x * 2^32 + w when `on` (one limb up, w and on uniform in the group); the limb that falls off the top is lost
template <int T, int L>
CG_FN void g_slide_up_if(const Grp<T> &g, u32 *x, u32 w, bool on) {
  u32 in = g_up<T>(g, x[L - 1], w);
  QF_UNROLL for (int i = L - 1; i > 0; i--) x[i] = on ? x[i - 1] : x[i];
  x[0] = on ? in : x[0];
}
// block of scratch -> registers and back (lane-blocked)
template <int T, int L> CG_FN void g_from_scr(const Grp<T> &g, u32 *x, const u32 *s) { QF_UNROLL for (int i = 0; i < L; i++) x[i] = s[g.lane * L + i]; }
template <int T, int L> CG_FN void g_to_scr(const Grp<T> &g, u32 *s, const u32 *x) { QF_UNROLL for (int i = 0; i < L; i++) s[g.lane * L + i] = x[i]; }

// ---- products by one limb and linear combinations ------------------------------------------------------------------
// r = a * w; returns the limb above the top (every lane)
template <int T, int L>
CG_FN u32 g_mul1(const Grp<T> &g, u32 *r, const u32 *a, u32 w) {
  u64 c = 0;
  QF_UNROLL for (int i = 0; i < L; i++) { c = (u64)a[i] * w + (c >> 32); r[i] = (u32)c; }
  i64 top = g_propagate<T, L, true>(g, r, (i64)(c >> 32));
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
// r = |x*A - y*B| with the sign known (neg: the combination is <= 0); x, y < 2^30.  Two multiply-add chains (one
// IMAD.WIDE per limb each) and one borrow chain; the three words that cross the lane boundary fit one signed word.
template <int T, int L>
CG_FN void g_lin_sub(const Grp<T> &g, u32 *r, u32 x, const u32 *A, u32 y, const u32 *B, bool neg) {
  u32 p = neg ? y : x, q = neg ? x : y;
  u32 U[L], V[L];
  u64 cu = 0, cv = 0;
  QF_UNROLL for (int i = 0; i < L; i++) {
    cu = (u64)p * (neg ? B[i] : A[i]) + (cu >> 32);
    cv = (u64)q * (neg ? A[i] : B[i]) + (cv >> 32);
    U[i] = (u32)cu;
    V[i] = (u32)cv;
  }
  u32 bo = 0;
  QF_UNROLL for (int i = 0; i < L; i++) {
    u64 d = (u64)U[i] - V[i] - bo;
    r[i] = (u32)d;
    bo = (u32)(d >> 63);
  }
  i64 cw = (i64)(cu >> 32) - (i64)(cv >> 32) - (i64)bo;
  g_propagate<T, L, false>(g, r, cw);
}
// r = x*A + y*B; returns the limb above the top (every lane).  x, y < 2^30
template <int T, int L>
CG_FN u32 g_lin_add(const Grp<T> &g, u32 *r, u32 x, const u32 *A, u32 y, const u32 *B) {
  u64 c = 0;
  QF_UNROLL for (int i = 0; i < L; i++) {
    c = (u64)x * A[i] + (u64)y * B[i] + (c >> 32);
    r[i] = (u32)c;
  }
  i64 top = g_propagate<T, L, false>(g, r, (i64)(c >> 32));
  return g_bcast<T>(g, (u32)top, T - 1);
}

// ---- division by one limb (lane-serial chains: the remainder / borrow travels from lane to lane) ------------------
// a mod w (w == 0 is answered as if w were 1)
template <int T, int L>
CG_FN u32 g_mod1(const Grp<T> &g, const u32 *a, u32 w) {
  w = w ? w : 1u;
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
// q = a / w exactly (w | a; w == 0 is answered as if w were 1)
template <int T, int L>
CG_FN void g_divexact1(const Grp<T> &g, u32 *q, const u32 *a, u32 w) {
  w = w ? w : 1u;
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
// out[0 .. K + nb) = A * B[0 .. nb)   (unsigned; nb a multiple of L and the same for every group of the warp; A: K limbs;
// out must not overlap A or B).  Sliding window: step k adds A * B[k] into an accumulator whose word 0 is then final (it
// is product limb k, stored by lane 0) and moves one limb down.
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
      u32 bk = B[k0 + r];
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

// Q[0 .. m) = NUM / D in Z / 2^(32 m) for odd D (Jebelean's exact division, least significant limb first); m a multiple
// of L, the same for every group of the warp.  NUM, D: K limbs.  Q must not overlap NUM or D.
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
// K-limb window; the groups of the warp step together, each for its own number of quotient limbs.  Returns false when
// D == 0 (R is then arbitrary).
template <int T, int L>
CG_NOINLINE bool g_mod_s(const Grp<T> g, const u32 *U, int nu, const u32 *D, u32 *R) {
  constexpr int K = T * L;
  static_assert(L >= 2, "two leading limbs are read from the top lane");
  g_sync<T>(g);
  // limb counts of D and U without lane-divergent scanning loops: every lane looks at its own blocks, the group votes
  int nd, nuu;
  {
    int td = 0, tu = 0;
    QF_UNROLL for (int i = 0; i < L; i++) { int p = g.lane * L + i; td = D[p] ? p + 1 : td; }
    QF_UNROLL for (int i = 0; i < 2 * L; i++) { int p = g.lane * 2 * L + i; tu = (p < nu && U[p]) ? p + 1 : tu; }
    u32 bd = g_ballot<T>(g, td != 0), bu = g_ballot<T>(g, tu != 0);
    int ld = bd ? 31 - clz32(bd) : 0, lu = bu ? 31 - clz32(bu) : 0;
    nd = (int)g_bcast<T>(g, (u32)td, ld);
    nuu = (int)g_bcast<T>(g, (u32)tu, lu);
  }
  nu = nuu;
  const bool ok = nd != 0;
  if (!ok) nd = 1;
  const int s = ok ? clz32(D[nd - 1]) : 0, ls = K - nd;
  // U' = U << (32 ls + s) has nup limbs; D' = D << (32 ls + s) fills K limbs with its top bit set
  const int nup = nu == 0 ? 0 : nu + ls + ((s && (U[nu - 1] >> (32 - s))) ? 1 : 0);
  const bool small = nup < K;  // U < D: the remainder is U
  const int np = small ? 0 : nup - K;
  u32 x[L], d[L];
  QF_UNROLL for (int i = 0; i < L; i++) {
    d[i] = sh_limb(D, nd, g.lane * L + i, ls, s);
    x[i] = small ? 0u : sh_limb(U, nu, np + g.lane * L + i, ls, s);
  }
  {
    int c = g_cmp<T, L>(g, x, d);
    g_sub_if<T, L>(g, x, x, d, !small && c >= 0);
  }
  const u32 dt = g_bcast<T>(g, d[L - 1], T - 1) | (ok ? 0u : 0x80000000u), v = recip32(dt | 0x80000000u);
  const int steps = (int)w_max<T>(g, (u32)np);
  QF_NOUNROLL for (int it = 0; it < steps; it++) {
    const int p = np - 1 - it;
    const bool on = p >= 0;
    u32 nxt = on ? sh_limb(U, nu, p, ls, s) : 0u;
    u32 rt = g_bcast<T>(g, x[L - 1], T - 1);
    u32 t0 = g_bcast<T>(g, x[L - 2], T - 1);
    g_slide_up_if<T, L>(g, x, nxt, on);
    // (rt : x) / d, with rt <= dt
    u32 rem;
    u32 qh = (rt >= dt) ? 0xffffffffu : div_2by1(rt, t0, dt | 0x80000000u, v, &rem);
    if (!on) qh = 0;
    u64 bo = g_submul1<T, L>(g, x, d, qh);
    bool neg = on && (u64)rt < bo;
    if (w_any<T>(g, neg)) {
      // went negative: add back (at most twice)
      u32 top = rt - (u32)bo;
      top += g_add_if<T, L>(g, x, x, d, neg);
      bool again = neg && top != 0;
      if (w_any<T>(g, again)) g_add_if<T, L>(g, x, x, d, again);
    }
  }
  // remainder = x >> (32 ls + s), or U itself
  u32 *tmp = scr_c<T, L>(g);
  g_sync<T>(g);
  g_to_scr<T, L>(g, tmp, x);
  g_sync<T>(g);
  QF_UNROLL for (int i = 0; i < L; i++) {
    int q = g.lane * L + i + ls;
    u32 lo = q < K ? tmp[q] : 0u, hi = q + 1 < K ? tmp[q + 1] : 0u;
    u32 lng = s ? ((lo >> s) | (hi << (32 - s))) : lo;
    int pu = g.lane * L + i;
    x[i] = small ? (pu < nu ? U[pu] : 0u) : lng;
  }
  g_to_scr<T, L>(g, R, x);
  g_sync<T>(g);
  return ok;
}

}  // namespace cofhe
