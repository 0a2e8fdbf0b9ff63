// cofhe_b200 -- scalar x ciphertext matmul as a batched multi-exponentiation (register-resident fast path).
//
//   out[i,k] = init o prod_j ct[i,j]^{s[j,k]}          ct: n x m ciphertexts, s: m x p non-negative scalars
//
// replaces CPUCryptoSystem::scal_ciphertext_tensors, matmul branch (reference
// include/x86_64/cpu_cryptosystem_tensor_ops.inl:346-461: qfi_nupow per (i,j) over the p scalars of row j, then the
// accumulation over j).  Only the reduced results are contractual, so the work is organised for the GPU instead:
//   1. k_mm_chain : per base form ct[i,j]:   B_w = ct[i,j]^(2^(W*w)),  w = 0..nwin-1      (W squarings per window)
//   2. k_mm_fill  : per (base, window):      T[w][d] = B_w^d,  d = 1..2^W-1               (one squaring, 2^W-3 comps)
//   3. k_mm_accum : per (chunk, i, k):       P = prod_{j in chunk} prod_w T_{ij}[w][digit_w(s[j,k])]   (comps only)
//   4. k_mm_final : per (i, k):              out = init o prod_chunks P
// Every product is a plain composition, so there is no squaring chain per output and the j range can be split
// into chunks freely to fill the machine when n*p is small.  Each base is reused by p outputs and each scalar by n.
// Elements the fast path declines anywhere are appended to the bail list and recomputed on the robust path.
#pragma once
#include "fast_ops.cuh"
#include "matmul_shape.cuh"
#if defined(__CUDACC__)
#include "fast_launch.cuh"
#endif

namespace cofhe {

// packed table entry stride (words), padded to 16 bytes
template <int N>
struct MMLayout {
  static constexpr int TE = ((3 * N + 1) + 3) & ~3;
};
// entry markers (a[0] == 0 never happens for a valid form whose a > 0 fits... use a full-zero `a` instead):
// valid form: a != 0.  a == 0 && b[0] == 0: "empty product".  a == 0 && b[0] == 1: "declined".
template <int N>
QF_HD void mm_mark(u32 *t, u32 kind) {
  QF_UNROLL for (int i = 0; i < MMLayout<N>::TE; i++) t[i] = 0;
  t[N] = kind;
}
template <int N>
QF_HD int mm_kind(const u32 *t) {  // 0 valid, 1 empty, 2 declined
  u32 o = 0;
  QF_UNROLL for (int i = 0; i < N; i++) o |= t[i];
  return o ? 0 : (t[N] ? 2 : 1);
}
template <int N>
QF_HD void mm_store(u32 *t, const FormR<N> &f) {
  fr_pack<N>(t, f);
  QF_UNROLL for (int i = 3 * N + 1; i < MMLayout<N>::TE; i++) t[i] = 0;
}

#if defined(__CUDACC__)

// 1. one thread per (y, base): B_w = ct^(2^(W*w)) into T[w][1]
template <int N, int CAP>
__global__ void __launch_bounds__(kMMBlock) k_mm_chain(const Group<CAP> *G, int dbits, int nc, MMShape s, const u32 *ct, u32 *tab) {
  constexpr int TE = MMLayout<N>::TE;
  if (gs_pick(dbits, blockIdx.y)) G += 1;  // compact variant: form block 0 (c1) lives in Cl(DeltaK)
  size_t nb = (size_t)s.n * s.m;
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool live = idx < nb;  // threads past the end idle inside the loop (block barriers)
  if (!live) idx = nb - 1;
  unsigned y = blockIdx.y;
  FormR<N> x;
  u32 xb[TE];
  bool ok = fr_load<N>(x, ct + (size_t)y * (2 * N + nc) * nb, nc, nb, idx);
  mm_store<N>(xb, x);
  int total = (s.nwin - 1) * s.W;
  for (int t = 0;; t++) {
    if (live && t % s.W == 0) {
      u32 *e = tab + mm_tab_index(s, y, idx, t / s.W, 1) * TE;
      if (ok) { QF_UNROLL for (int i = 0; i < TE; i++) e[i] = xb[i]; }
      else mm_mark<N>(e, 1);
    }
    if (t >= total) break;
    QF_BLOCK_ALIGN();
    if (ok && live) ok = cc_compose<N, CAP, true>(G, xb, xb, false, dbits);
  }
}

// 2. one thread per (y, base, window): T[w][d] = B_w^d
template <int N, int CAP>
__global__ void __launch_bounds__(kMMBlock) k_mm_fill(const Group<CAP> *G, int dbits, MMShape s, u32 *tab) {
  constexpr int TE = MMLayout<N>::TE;
  if (gs_pick(dbits, blockIdx.y)) G += 1;
  size_t total = (size_t)s.n * s.m * s.nwin;
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool live = idx < total;
  if (!live) idx = total - 1;
  unsigned y = blockIdx.y;
  size_t base = idx / s.nwin;
  int win = (int)(idx % s.nwin);
  u32 *e = tab + mm_tab_index(s, y, base, win, 1) * TE;
  u32 xb[TE];
  bool ok = mm_kind<N>(e) == 0;
  QF_UNROLL for (int i = 0; i < TE; i++) xb[i] = e[i];
  int nt = (1 << s.W) - 1;
  for (int d = 2; d <= nt; d++) {
    QF_BLOCK_ALIGN();
    if (ok && live) ok = d == 2 ? cc_compose<N, CAP, true>(G, xb, xb, false, dbits) : cc_compose<N, CAP, false>(G, xb, e, false, dbits);
    if (live) {
      u32 *o = e + (size_t)(d - 1) * TE;
      if (ok) { QF_UNROLL for (int i = 0; i < TE; i++) o[i] = xb[i]; }
      else mm_mark<N>(o, 1);
    }
  }
}

// 3. one thread per (chunk, y, i, k): product over the chunk's j range and all windows; compositions only
template <int N, int CAP>
__global__ void __launch_bounds__(kMMBlock) k_mm_accum(const Group<CAP> *G, int dbits, MMShape s, const u32 *scal, const u32 *tab, u32 *part) {
  constexpr int TE = MMLayout<N>::TE;
  if (gs_pick(dbits, blockIdx.y)) G += 1;
  size_t no = (size_t)s.n * s.p;
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool live = idx < no;
  if (!live) idx = no - 1;
  unsigned y = blockIdx.y, c = blockIdx.z;
  unsigned i = (unsigned)(idx / s.p), k = (unsigned)(idx % s.p);
  unsigned j0 = c * s.jper, j1 = j0 + s.jper < s.m ? j0 + s.jper : s.m;
  size_t sstride = (size_t)s.m * s.p;
  u32 xb[TE];
  bool ok = true, started = false;
  // the trip counts are the same for every thread of the block (block barriers inside)
  for (unsigned j = j0; j < j1; j++) {
    const u32 *e = scal + (size_t)j * s.p + k;
    for (int w = 0; w < s.nwin; w++) {
      u32 d = (ok && live) ? exp_digit(e, s.el, sstride, w, s.W) : 0u;
      const u32 *te = tab + mm_tab_index(s, y, (size_t)i * s.m + j, w, d ? (int)d : 1) * TE;
      if (d && mm_kind<N>(te) != 0) { ok = false; d = 0; }
      // every thread of the block makes the call (barriers inside); threads without a digit here just walk through
      const bool act = d != 0 && started;
      if (QF_BLOCK_ANY(act)) { bool k = cc_compose<N, CAP, false, true>(G, xb, te, false, dbits, act); if (act) ok = k; }
      if (d && !started) {
        QF_UNROLL for (int i = 0; i < TE; i++) xb[i] = te[i];
        started = true;
      }
    }
  }
  if (live) {
    u32 *o = part + (((size_t)c * 2 + y) * no + idx) * TE;
    if (!ok) mm_mark<N>(o, 1);
    else if (!started) mm_mark<N>(o, 0);
    else { QF_UNROLL for (int i = 0; i < TE; i++) o[i] = xb[i]; }
  }
}

// 4. one thread per (y, i, k): out = init o prod_chunks part   (init == nullptr: plain product)
template <int N, int CAP>
__global__ void __launch_bounds__(kMMBlock) k_mm_final(const Group<CAP> *G, int dbits, int nc, MMShape s, const u32 *init, const u32 *part, u32 *out,
                                                       u32 *bail, int *nbail) {
  constexpr int TE = MMLayout<N>::TE;
  if (gs_pick(dbits, blockIdx.y)) G += 1;
  size_t no = (size_t)s.n * s.p;
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  bool live = idx < no;
  if (!live) idx = no - 1;
  unsigned y = blockIdx.y;
  FormR<N> x;
  u32 xb[TE];
  bool started = init != nullptr;
  bool ok = started ? fr_load<N>(x, init + (size_t)y * (2 * N + nc), nc, 1, 0) : true;
  if (started) mm_store<N>(xb, x);
  for (unsigned c = 0; c < s.chunks; c++) {
    const u32 *pe = part + (((size_t)c * 2 + y) * no + idx) * TE;
    int kind = (ok && live) ? mm_kind<N>(pe) : 1;
    if (kind == 2) ok = false;
    QF_BLOCK_ALIGN();
    if (kind == 0) {
      if (started) ok = cc_compose<N, CAP, false>(G, xb, pe, false, dbits);
      else {
        QF_UNROLL for (int i = 0; i < TE; i++) xb[i] = pe[i];
        started = true;
      }
    }
  }
  ok = ok && started;  // an empty product without init is the neutral form: robust path
  if (live) {
    if (ok) {
      fr_unpack<N>(x, xb);
      fr_store<N>(x, out + (size_t)y * (2 * N + nc) * no, nc, no, idx);
    } else bail[atomicAdd(nbail, 1)] = (u32)(y * no + idx);
  }
}

// robust path: all outputs (list == nullptr) or the listed ones; clears the next call's counter
template <int CAP>
__global__ void __launch_bounds__(128) k_mm_robust(const Group<CAP> *G, Shape sh, MMShape s, const u32 *ct, const u32 *scal,
                                                        const u32 *init, u32 *out, const u32 *bail, const int *nbail,
                                                        int *nbail_next, int *err, int alt) {
  size_t no = (size_t)s.n * s.p, nb = (size_t)s.n * s.m;
  size_t cnt = bail ? (size_t)*nbail : 2 * no;
  if (nbail_next && blockIdx.x == 0 && threadIdx.x == 0) *nbail_next = 0;
  for (size_t q = (size_t)blockIdx.x * blockDim.x + threadIdx.x; q < cnt; q += (size_t)gridDim.x * blockDim.x) {
    size_t e = bail ? bail[q] : q;
    size_t y = e / no, idx = e % no;
    unsigned i = (unsigned)(idx / s.p), k = (unsigned)(idx % s.p);
    Err er; er.code = 0;
    Form<CAP> acc, f, t;
    const Group<CAP> &Gy = G[y == 0 ? alt : 0];  // compact variant: form block 0 (c1) lives in Cl(DeltaK)
    if (init) form_load(er, acc, init + y * form_limbs(sh), sh, 1, 0);
    else qf_identity(er, Gy, acc);
    for (unsigned j = 0; j < s.m && !er.code; j++) {
      const u32 *ex = scal + (size_t)j * s.p + k;
      bool zero = true;
      for (int l = 0; l < s.el; l++) zero = zero && ex[(size_t)l * s.m * s.p] == 0;
      if (zero) continue;
      form_load(er, f, ct + y * form_limbs(sh) * nb, sh, nb, (size_t)i * s.m + j);
      qf_pow(er, Gy, t, f, ex, s.el, (size_t)s.m * s.p, false);
      qf_compose(er, Gy, acc, acc, t, false, false);
    }
    form_store(er, acc, out + y * form_limbs(sh) * no, sh, no, idx);
    if (er.code) atomicMax(err, er.code);
  }
}
#endif

}  // namespace cofhe
