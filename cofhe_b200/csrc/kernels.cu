// cofhe_b200 -- CUDA kernels and the C ABI (include/cofhe_b200.h).  sm_100a only.
//
// One thread owns one form operation (independent elements of a batch).  Buffers are
// structure-of-arrays limbs so that a warp's loads/stores of one limb row are one coalesced 128-byte line.
#include <cuda_runtime.h>
#include <stdio.h>
#include <stdlib.h>
#include <string.h>
#include <time.h>
#include <mutex>
#include <vector>
#include <new>
#include <utility>

#include "cofhe_b200.h"
#include "coop_launch.cuh"
#include "codec.cuh"
#include "fast_launch.cuh"
#include "fast_ops.cuh"
#include "matmul.cuh"
#include "scheme_ops.cuh"
#include "wire.cuh"

using namespace cofhe;

struct cofhe_b200_ctx {
  int device;
  int sms;            // multiprocessors of the device (grid sizing)
  int cap;            // capacity class of the generic integer layer
  u32 na, nc, nd, dbits;
  u32 h_absd[2 * 216];  // |Delta| (host copy)
  void *d_group;      // Group<CAP>[2] in device memory: the group of Delta, then the group of DeltaK (compact variant)
  // compact variant (cofhe_b200_set_compact): c1 halves, h and pk live in Cl(DeltaK); stored in the same limb rows
  int compact;
  u32 dbitsK;
  void *d_groupK;     // = &((Group<CAP>*)d_group)[1]
  int gforce;         // > 0: the launches issued right now run entirely in Cl(DeltaK) (UseK)
  unsigned char *d_lift = nullptr;  // chi flags of cofhe_b200_lift, one byte per element
  size_t lift_bytes = 0;
  int *d_err;         // sticky per-context element status (max of Err codes)
  cudaStream_t xfer;  // internal stream of the _host entry points
  uint64_t launches;
  int fastn;          // limbs of the register-resident instantiation serving this shape (0: none, robust path only)
  int coopn;          // limbs of the cooperative (lane-group) instantiation serving this shape (0: none)
  u32 *d_bail;        // elements the fast path declined, redone by the sparse robust kernels
  int *d_nbail;       // two counters used alternately (the sparse kernel of call k clears the one of call k+1)
  size_t bail_cap;
  int flip;
  // pipeline lanes of the _host entry points: own stream and own bail list, so that the chunks of one call overlap
  // (H2D of chunk i+1 | kernels of chunk i | D2H of chunk i-1)
  struct Lane { cudaStream_t st; u32 *d_bail; int *d_nbail; size_t bail_cap; int flip; } lane[4];
  // staging of the _host entry points (grown on demand, reused across calls)
  u32 *d_stage;
  size_t stage_bytes;
  // scratch of the composite operations (matmul tables and partial products, fixed-base digits, ...)
  u32 *d_work;
  size_t work_bytes;
  u32 *d_powtab = nullptr;   // window tables of the cooperative exponentiation kernel
  size_t powtab_bytes = 0;
  u32 *d_tmp;         // form / ciphertext temporaries of encrypt, combine, decrypt
  size_t tmp_bytes;
  // CL_HSM2k scheme (set_scheme) and fixed bases h, pk (set_bases)
  void *d_scheme;     // Scheme<CAP> in device memory
  int k, kl;
  u32 *d_exp;         // one shared exponent (share / secret key) staged from host memory: 64 limbs
  u32 *d_fb_bases;    // (h, pk) as one ciphertext-shaped element [2][fl][1]
  u32 *d_fb_tab;      // window tables of h and pk (fast tiers)
  MMShape fb;
  int fb_set;
  int fb_coop = 0;    // the fixed-base tables are SoA batches for the cooperative multi-exponentiation (coop_tables)
  size_t fb_ne = 0;   // their entries per form block
  u32 *d_tpbail = nullptr;  // outputs the cooperative multi-exponentiation hands to the robust path: [0] counter, [1] spare, [2..] list
  size_t tpbail_cap = 0;
  // One context may be shared by threads and streams (include/cofhe_b200.h "Threads and streams"): entry points
  // serialise on `mu`; when the caller's stream changes, the new stream is ordered after the previous one, because the
  // scratch above is shared by all calls.
  std::recursive_mutex mu;
  cudaStream_t last;
  bool busy;
  cudaEvent_t order;
  // tuning knobs (cofhe_b200_set_option; the COFHE_B200_* environment variables are read once, at creation)
  // device blocks handed out by cofhe_b200_dev_alloc (live) and returned ones kept for reuse (pool)
  struct Block { void *p; size_t bytes; };
  std::vector<Block> live, pool;
  size_t pool_bytes = 0;
  struct Opt { long no_fast, mm_w, mm_chunks, mm_block, fb_w, host_staged, host_loadc, host_chunks, host_trace, coop, coop_mb, coop_pb, fast_block, fast_flat; } opt;
};

namespace {


#define CUDA_TRY(x)                                   \
  do {                                                \
    cudaError_t _e = (x);                             \
    if (_e != cudaSuccess) return COFHE_B200_ECUDA;   \
  } while (0)

#define CAP_DISPATCH(cap, ...)                                   \
  switch (cap) {                                                 \
    case 32: { constexpr int CAP = 32; __VA_ARGS__; } break;     \
    case 48: { constexpr int CAP = 48; __VA_ARGS__; } break;     \
    case 136: { constexpr int CAP = 136; __VA_ARGS__; } break;   \
    case 216: { constexpr int CAP = 216; __VA_ARGS__; } break;   \
    default: return COFHE_B200_EINVAL;                           \
  }

int cap_for(u32 nd) {
  u32 need = 2 * nd + 4;
  const int caps[] = {32, 48, 136, 216};
  for (int c : caps)
    if ((u32)c >= need) return c;
  return 0;
}

inline unsigned grid_for(size_t n) { return (unsigned)((n + kBlock - 1) / kBlock); }

// grid.y selects the form block inside a ciphertext batch (0: c1, 1: c2); plain form batches use grid.y == 1
template <int CAP>
__global__ void __launch_bounds__(kBlock) k_compose(const Group<CAP> *G, Shape s, size_t n, const u32 *in1, const u32 *in2,
                                                    u32 *out, int neg2, int *err, int alt) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  size_t off = (size_t)blockIdx.y * form_limbs(s) * n;
  int c = op_compose<CAP>(G[blockIdx.y == 0 ? alt : 0], s, n, idx, in1 + off, in2 + off, out + off, neg2 != 0);
  if (c) atomicMax(err, c);
}
template <int CAP>
__global__ void __launch_bounds__(kBlock) k_square(const Group<CAP> *G, Shape s, size_t n, const u32 *in, u32 *out, int *err, int alt) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  size_t off = (size_t)blockIdx.y * form_limbs(s) * n;
  int c = op_square<CAP>(G[blockIdx.y == 0 ? alt : 0], s, n, idx, in + off, out + off);
  if (c) atomicMax(err, c);
}
template <int CAP>
__global__ void __launch_bounds__(kBlock) k_reduce(const Group<CAP> *G, Shape sin, Shape sout, size_t n, const u32 *in, u32 *out,
                                                   int *err) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  int c = op_reduce<CAP>(*G, sin, sout, n, idx, in, out);
  if (c) atomicMax(err, c);
}
template <int CAP>
__global__ void __launch_bounds__(kBlock) k_pow(const Group<CAP> *G, Shape s, size_t n, const u32 *in, u32 *out, const u32 *exp,
                                                int el, int per_elem, const unsigned char *expneg, int shared_neg, int *err, int alt) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  size_t off = (size_t)blockIdx.y * form_limbs(s) * n;
  int c = op_pow<CAP>(G[blockIdx.y == 0 ? alt : 0], s, n, idx, in + off, out + off, exp, el, per_elem != 0, expneg, shared_neg != 0);
  if (c) atomicMax(err, c);
}


// ---- sparse robust redo of what the register-resident kernels (fast_launch.cuh) decline ---------------------------
// op: 0 compose, 1 square, 2 pow.  Grid-stride over the bail list; clears the counter of the next call.
template <int CAP>
__global__ void __launch_bounds__(kBlock) k_redo(const Group<CAP> *G, Shape s, size_t n, int op, const u32 *in1, const u32 *in2,
                                                 u32 *out, int neg2, const u32 *exp, int el, int per_elem,
                                                 const unsigned char *expneg, int shared_neg, const u32 *bail, const int *nbail,
                                                 int *nbail_next, int *err, int alt) {
  int cnt = *nbail;
  if (blockIdx.x == 0 && threadIdx.x == 0) *nbail_next = 0;
  for (int i = blockIdx.x * blockDim.x + threadIdx.x; i < cnt; i += gridDim.x * blockDim.x) {
    u32 e = bail[i];
    size_t y = e / n, idx = e % n, off = y * form_limbs(s) * n;
    int c;
    const Group<CAP> &Gy = G[y == 0 ? alt : 0];
    if (op == 0) c = op_compose<CAP>(Gy, s, n, idx, in1 + off, in2 + off, out + off, (neg2 & 1) != 0, (neg2 & 2) != 0);
    else if (op == 1) c = op_square<CAP>(Gy, s, n, idx, in1 + off, out + off);
    else c = op_pow<CAP>(Gy, s, n, idx, in1 + off, out + off, exp, el, per_elem != 0, expneg, shared_neg != 0);
    if (c) atomicMax(err, c);
  }
}

// (limbs of a and b, capacity class of the robust layer) pairs that exist: cap_for(nd) with nd ~ 2*na - 1
#ifdef COFHE_B200_DEV_N7  /* development builds: only the T0 instantiation (compiles in a quarter of the time) */
#define FAST_DISPATCH(fn, ...)                                                        \
  switch (fn) {                                                                       \
    case 7: { constexpr int FN = 7, FCAP = 32; __VA_ARGS__; } break;                  \
    default: return COFHE_B200_EINVAL;                                                \
  }
#define FAST_HAVE(na) ((na) == 7)
#else
#define FAST_DISPATCH(fn, ...)                                                        \
  switch (fn) {                                                                       \
    case 2: { constexpr int FN = 2, FCAP = 32; __VA_ARGS__; } break;                  \
    case 4: { constexpr int FN = 4, FCAP = 32; __VA_ARGS__; } break;                  \
    case 7: { constexpr int FN = 7, FCAP = 32; __VA_ARGS__; } break;                  \
    case 11: { constexpr int FN = 11, FCAP = 48; __VA_ARGS__; } break;                \
    default: return COFHE_B200_EINVAL;                                                \
  }
#define FAST_HAVE(na) ((na) == 2 || (na) == 4 || (na) == 7 || (na) == 11)
#endif
int fast_cap(int fn) { return fn == 11 ? 48 : 32; }
int fast_n_for(u32 na, int cap, bool no_fast) {
  bool have = FAST_HAVE(na) && fast_cap((int)na) == cap;
  return have && !no_fast ? (int)na : 0;
}
long env_opt(const char *name, long lo, long hi) {
  const char *v = getenv(name);
  if (!v) return -1;
  long x = atol(v);
  return x >= lo && x <= hi ? x : -1;
}

// cooperative instantiations: (lanes per form, limbs per lane) per coefficient width; coop_n<N>.cu
#ifdef COFHE_B200_DEV_N7
#define COOP_DISPATCH(cn, ...)                                                                   \
  switch (cn) {                                                                                  \
    case 7: { constexpr int CT = 2, CL = 5, CN = 7, CCAP = 32; __VA_ARGS__; } break;             \
    default: return COFHE_B200_EINVAL;                                                           \
  }
#define COOP_HAVE(na) ((na) == 7)
#else
#define COOP_DISPATCH(cn, ...)                                                                   \
  switch (cn) {                                                                                  \
    case 7: { constexpr int CT = 2, CL = 5, CN = 7, CCAP = 32; __VA_ARGS__; } break;             \
    case 11: { constexpr int CT = 2, CL = 7, CN = 11, CCAP = 48; __VA_ARGS__; } break;           \
    case 33: { constexpr int CT = 4, CL = 9, CN = 33, CCAP = 136; __VA_ARGS__; } break;          \
    case 53: { constexpr int CT = 8, CL = 7, CN = 53, CCAP = 216; __VA_ARGS__; } break;          \
    default: return COFHE_B200_EINVAL;                                                           \
  }
#define COOP_HAVE(na) ((na) == 7 || (na) == 11 || (na) == 33 || (na) == 53)
#endif
int coop_cap(int cn) { return cn == 53 ? 216 : (cn == 33 ? 136 : (cn == 11 ? 48 : 32)); }
// which path serves a batched form operation: the cooperative kernels where no register instantiation exists (the large
// discriminants), or when forced ("coop" option: 1 always, 0 never)
bool use_coop(const cofhe_b200_ctx *ctx) {
  if (!ctx->coopn || ctx->opt.coop == 0) return false;
  return ctx->opt.coop == 1 || !ctx->fastn;
}

// Which class group serves the form blocks of a launch (compact variant: bn.cuh gs_pick).  G / dbits: what the register
// and cooperative kernels receive; alt: what the generic kernels receive (index of the group serving form block 0).
struct GSel { const void *G; int dbits; int alt; };
GSel gsel(const cofhe_b200_ctx *ctx, unsigned ny) {
  if (ctx->compact && ctx->gforce) return GSel{ctx->d_groupK, (int)ctx->dbitsK, 0};
  if (ctx->compact && ny == 2) return GSel{ctx->d_group, (int)(ctx->dbits | (ctx->dbitsK << 16)), 1};
  return GSel{ctx->d_group, (int)ctx->dbits, 0};
}
// the launches issued while one of these lives run entirely in Cl(DeltaK) (no effect outside the compact variant)
struct UseK {
  cofhe_b200_ctx *c;
  explicit UseK(cofhe_b200_ctx *ctx) : c(ctx) { c->gforce++; }
  ~UseK() { c->gforce--; }
};

int ensure_bail(cofhe_b200_ctx *ctx, size_t entries) {
  if (entries >> 32) return COFHE_B200_EINVAL;
  if (entries <= ctx->bail_cap) return COFHE_B200_OK;
  if (ctx->d_bail) cudaFree(ctx->d_bail);  // synchronises with kernels still reading the old list
  ctx->d_bail = nullptr; ctx->bail_cap = 0;
  if (cudaMalloc(&ctx->d_bail, entries * sizeof(u32)) != cudaSuccess) return COFHE_B200_ENOMEM;
  ctx->bail_cap = entries;
  return COFHE_B200_OK;
}

// ---- scheme-level kernels (scheme_ops.cuh) ----------------------------------------------------------------------
template <int KL, int CAP>
__global__ void __launch_bounds__(kBlock) k_power_of_f(const Group<CAP> *G, const Scheme<CAP> *S, Shape s, size_t n, const u32 *m,
                                                       int ml, u32 *out, int *err) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  int c = op_power_of_f<KL, CAP>(*G, *S, s, n, idx, m, ml, out);
  if (c) atomicMax(err, c);
}
template <int KL, int CAP>
__global__ void __launch_bounds__(kBlock) k_dlog(const Group<CAP> *G, const Scheme<CAP> *S, Shape s, size_t n, const u32 *in,
                                                 u32 *out, int ml, int *err) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  int c = op_dlog<KL, CAP>(*G, *S, s, n, idx, in, out, ml);
  if (c) atomicMax(err, c);
}
// ---- compact variant: Cl(DeltaK) -> Cl(Delta) (CL_HSM2k::from_Cl_DeltaK_to_Cl_Delta, CL_HSM2k.inl:458-474): step 1 and
// step 3 as kernels around scheme_ops.cuh's per-element bodies; step 2 (f^(2^k)) runs on the exponentiation kernels
template <int CAP>
__global__ void __launch_bounds__(kBlock) k_lift_2exp(Shape s, size_t n, int k, const u32 *in, u32 *out, unsigned char *chi, int *err) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  int c = op_lift_2exp<CAP>(s, n, idx, k, in, out, chi);
  if (c) atomicMax(err, c);
}
template <int CAP>
__global__ void __launch_bounds__(kBlock) k_lift_fix(const Group<CAP> *G, const Scheme<CAP> *S, Shape s, size_t n, u32 *io,
                                                     const unsigned char *chi, int *err) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  if (!chi[idx]) return;
  int c = op_lift_fix<CAP>(*G, *S, s, n, idx, io);
  if (c) atomicMax(err, c);
}

// out[row][idx] = in[row][0] for `rows` limb rows
__global__ void k_bcast(size_t n, int rows, const u32 *in, u32 *out) {
  size_t idx = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (idx >= n) return;
  for (int r = 0; r < rows; r++) out[(size_t)r * n + idx] = in[r];
}

#define SCHEME_DISPATCH(ctx, ...)                                                                   \
  CAP_DISPATCH((ctx)->cap, switch ((ctx)->kl) {                                                       \
    case 1: { constexpr int KL = 1; __VA_ARGS__; } break;                                            \
    case 4: { constexpr int KL = 4; __VA_ARGS__; } break;                                            \
    case 8: { constexpr int KL = 8; __VA_ARGS__; } break;                                            \
    default: return COFHE_B200_ESTATE;                                                              \
  })

// out[.][e] = a[.][e] * b[.][e] mod 2^(32 ml)  (plaintexts are used modulo 2^k only: f has order 2^k)
__global__ void __launch_bounds__(kBlock) k_mul_lo(size_t n, int ml, const u32 *a, const u32 *b, u32 *out) {
  size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  u32 r[16];
  for (int i = 0; i < ml; i++) r[i] = 0;
  for (int i = 0; i < ml; i++) {
    unsigned long long c = 0;
    u32 ai = a[(size_t)i * n + e];
    for (int j = 0; i + j < ml; j++) {
      c += (unsigned long long)ai * b[(size_t)j * n + e] + r[i + j];
      r[i + j] = (u32)c;
      c >>= 32;
    }
  }
  for (int i = 0; i < ml; i++) out[(size_t)i * n + e] = r[i];
}

int ensure_tmp(cofhe_b200_ctx *ctx, size_t bytes) {
  if (bytes <= ctx->tmp_bytes) return COFHE_B200_OK;
  if (ctx->d_tmp) cudaFree(ctx->d_tmp);
  ctx->d_tmp = nullptr; ctx->tmp_bytes = 0;
  if (cudaMalloc(&ctx->d_tmp, bytes) != cudaSuccess) { cudaGetLastError(); return COFHE_B200_ENOMEM; }
  ctx->tmp_bytes = bytes;
  return COFHE_B200_OK;
}

int ensure_work(cofhe_b200_ctx *ctx, size_t bytes) {
  if (bytes <= ctx->work_bytes) return COFHE_B200_OK;
  if (ctx->d_work) cudaFree(ctx->d_work);  // synchronises with kernels still using the old scratch
  ctx->d_work = nullptr; ctx->work_bytes = 0;
  if (cudaMalloc(&ctx->d_work, bytes) != cudaSuccess) { cudaGetLastError(); return COFHE_B200_ENOMEM; }
  ctx->work_bytes = bytes;
  return COFHE_B200_OK;
}

int check_launch(cofhe_b200_ctx *ctx) {
  ctx->launches++;
  return cudaGetLastError() == cudaSuccess ? COFHE_B200_OK : COFHE_B200_ECUDA;
}


// One batched form operation (op: 0 compose, 1 square, 2 pow) over ny form blocks of n elements.
int launch_op(cofhe_b200_ctx *ctx, int op, unsigned ny, size_t n, const u32 *in1, const u32 *in2, u32 *out, int neg2,
              const u32 *exp, int el, int per_elem, const unsigned char *expneg, int shared_neg, cudaStream_t st) {
  Shape s{(int)ctx->na, (int)ctx->nc};
  dim3 grid(grid_for(n), ny);
  const bool coop = use_coop(ctx) && !(neg2 & 2);
  const GSel gs = gsel(ctx, ny);
  if (!ctx->fastn && !coop) {
    CAP_DISPATCH(ctx->cap, {
      const Group<CAP> *G = (const Group<CAP> *)gs.G;
      if (op == 0) k_compose<CAP><<<grid, kBlock, 0, st>>>(G, s, n, in1, in2, out, neg2 & 1, ctx->d_err, gs.alt);
      else if (op == 1) k_square<CAP><<<grid, kBlock, 0, st>>>(G, s, n, in1, out, ctx->d_err, gs.alt);
      else k_pow<CAP><<<grid, kBlock, 0, st>>>(G, s, n, in1, out, exp, el, per_elem, expneg, shared_neg, ctx->d_err, gs.alt);
    })
    return check_launch(ctx);
  }
  int rc = ensure_bail(ctx, (size_t)ny * n);
  if (rc != COFHE_B200_OK) return rc;
  int *cnt = ctx->d_nbail + ctx->flip, *cnt_next = ctx->d_nbail + (ctx->flip ^ 1);
  ctx->flip ^= 1;
  int dbits = gs.dbits, nc = (int)ctx->nc;
  if (coop) {
    COOP_DISPATCH(ctx->coopn, {
      typedef CoopKernels<CT, CL, CN, CCAP> K;
      // default: block barriers between the phases; 256-thread blocks once the batch fills the machine several times over
      if (op < 2) K::compose(op == 1, ny, st, gs.G, dbits, nc, n, in1, in2, out, neg2, ctx->d_bail, cnt,
                             ctx->opt.coop_mb > 0 ? (int)ctx->opt.coop_mb : ((size_t)ny * n >= 98304 ? 24 : 14));
      else {
        // 256-thread blocks once every SM gets one (measured at T1: 16 384 exponentiations 690 -> 521 ms, 65 536: 2.06 -> 1.79 s;
        // 2 048: 301 -> 353 ms)
        const size_t pow_lanes = ctx->coopn >= 53 ? 8 : (ctx->coopn >= 33 ? 4 : 2);
        const int pow_block = ctx->opt.coop_pb == 256 ? 256 : (ctx->opt.coop_pb == 4 ? 4 : (ctx->opt.coop_pb == 128 ? 128 :
                              ((size_t)ny * n >= (size_t)(ctx->sms > 0 ? ctx->sms : 148) * (256 / pow_lanes) ? 256 : 128)));
        size_t tw = K::pow_tab_words(ny, nc, n) * sizeof(u32);
        if (tw > ctx->powtab_bytes) {
          if (ctx->d_powtab) cudaFree(ctx->d_powtab);
          ctx->d_powtab = nullptr; ctx->powtab_bytes = 0;
          if (cudaMalloc(&ctx->d_powtab, tw) != cudaSuccess) { cudaGetLastError(); return COFHE_B200_ENOMEM; }
          ctx->powtab_bytes = tw;
        }
        K::pow(ny, st, gs.G, dbits, nc, n, in1, out, exp, el, per_elem, expneg, shared_neg, ctx->d_powtab, ctx->d_bail, cnt, pow_block);
      }
    })
  } else {
    const int pb = pick_block((size_t)ny * n, ctx->sms);
    FAST_DISPATCH(ctx->fastn, {
      typedef FastKernels<FN, FCAP> K;
      // option "fast_block" (256 / 512): wider blocks for the composition kernel (tools/quick_time.py measures them)
      // measured at T0 (profiles/r2n_quick_time.txt): 256-thread blocks +6-7 % from 2^18 compositions, 512 slower everywhere
      const int wide = (ctx->opt.fast_block == 256 || ctx->opt.fast_block == 512) ? (int)ctx->opt.fast_block
                       : (ctx->opt.fast_block == 128 ? 0 : ((size_t)ny * n >= 200000 ? 256 : 0));
      // branch-free scalar Euclid rounds: -5 % on launches that last as long as one warp (up to ~1.5 warps per scheduler),
      // +3-6 % once the machine is full (profiles/r2o_quick_time.txt); automatic by batch size unless the option says otherwise
      const bool flat = ctx->opt.fast_flat == 1 || (ctx->opt.fast_flat < 0 && !wide && (size_t)ny * n <= 49152);
      if (op < 2) K::compose(op == 1, grid, st, gs.G, dbits, nc, n, in1, in2, out, neg2 | (wide << 8) | (flat ? (1 << 20) : 0), ctx->d_bail, cnt);
      else K::pow(dim3((unsigned)((n + pb - 1) / pb), ny), pb, st, gs.G, dbits, nc, n, in1, out, exp, el, per_elem, expneg,
                  shared_neg, ctx->d_bail, cnt);
    })
  }
  rc = check_launch(ctx);
  if (rc != COFHE_B200_OK) return rc;
  unsigned rb = (unsigned)((size_t)ny * n < 4096 ? ((size_t)ny * n + kBlock - 1) / kBlock : 32);
  CAP_DISPATCH(ctx->cap, k_redo<CAP><<<rb, kBlock, 0, st>>>((const Group<CAP> *)gs.G, s, n, op, in1, in2, out, neg2, exp, el,
                                                          per_elem, expneg, shared_neg, ctx->d_bail, cnt, cnt_next, ctx->d_err, gs.alt))
  return check_launch(ctx);
}

// ---- matmul / fixed-base machinery (matmul.cuh) ---------------------------------------------------------------
size_t mm_tab_words(const cofhe_b200_ctx *ctx, const MMShape &s) {
  int te = ((3 * ctx->fastn + 1) + 3) & ~3;
  return (size_t)2 * s.n * s.m * s.nwin * ((1u << s.W) - 1) * te;
}
size_t mm_part_words(const cofhe_b200_ctx *ctx, const MMShape &s) {
  int te = ((3 * ctx->fastn + 1) + 3) & ~3;
  return (size_t)s.chunks * 2 * s.n * s.p * te;
}
// window tables of the n*m base ciphertexts `ct` into tab (fast tiers only)
int mm_tables(cofhe_b200_ctx *ctx, const MMShape &s, const u32 *ct, u32 *tab, cudaStream_t st) {
  const GSel gs = gsel(ctx, 2);
  int dbits = gs.dbits, nc = (int)ctx->nc;
  size_t nb = (size_t)s.n * s.m;
  FAST_DISPATCH(ctx->fastn, {
    typedef FastKernels<FN, FCAP> K;
    const int b1 = pick_block(2 * nb, ctx->sms), b2 = pick_block(2 * nb * s.nwin, ctx->sms);
    K::mm_chain(dim3((unsigned)((nb + b1 - 1) / b1), 2), b1, st, gs.G, dbits, nc, s, ct, tab);
    ctx->launches++;
    K::mm_fill(dim3((unsigned)((nb * s.nwin + b2 - 1) / b2), 2), b2, st, gs.G, dbits, s, tab);
  })
  return check_launch(ctx);
}
// out[i,k] = init o prod_j prod_w tab_{ij}[w][digit_w(scal[j,k])]; declined outputs are redone from ct on the robust path
int mm_eval(cofhe_b200_ctx *ctx, const MMShape &s, const u32 *ct, const u32 *scal, const u32 *init, u32 *out, const u32 *tab,
            u32 *part, cudaStream_t st) {
  Shape sh{(int)ctx->na, (int)ctx->nc};
  size_t no = (size_t)s.n * s.p;
  int rc = ensure_bail(ctx, 2 * no);
  if (rc != COFHE_B200_OK) return rc;
  const GSel gs = gsel(ctx, 2);
  int dbits = gs.dbits, nc = (int)ctx->nc;
  int *cnt = ctx->d_nbail + ctx->flip, *cnt_next = ctx->d_nbail + (ctx->flip ^ 1);
  ctx->flip ^= 1;
  FAST_DISPATCH(ctx->fastn, {
    typedef FastKernels<FN, FCAP> K;
    const int ba = s.block ? s.block : pick_block(2 * no * s.chunks, ctx->sms), bf = pick_block(2 * no, ctx->sms);
    K::mm_accum(dim3((unsigned)((no + ba - 1) / ba), 2, s.chunks), ba, st, gs.G, dbits, s, scal, tab, part);
    ctx->launches++;
    K::mm_final(dim3((unsigned)((no + bf - 1) / bf), 2), bf, st, gs.G, dbits, nc, s, init, part, out, ctx->d_bail, cnt);
    ctx->launches++;
    k_mm_robust<FCAP><<<32, 128, 0, st>>>((const Group<FCAP> *)gs.G, sh, s, ct, scal, init, out, ctx->d_bail, cnt, cnt_next, ctx->d_err, gs.alt);
  })
  return check_launch(ctx);
}
int mm_robust_all(cofhe_b200_ctx *ctx, const MMShape &s, const u32 *ct, const u32 *scal, const u32 *init, u32 *out, cudaStream_t st) {
  Shape sh{(int)ctx->na, (int)ctx->nc};
  size_t no = (size_t)s.n * s.p;
  unsigned blocks = (unsigned)((2 * no + 127) / 128);
  const GSel gs = gsel(ctx, 2);
  CAP_DISPATCH(ctx->cap, k_mm_robust<CAP><<<blocks, 128, 0, st>>>((const Group<CAP> *)gs.G, sh, s, ct, scal, init, out,
                                                                     nullptr, nullptr, nullptr, ctx->d_err, gs.alt))
  return check_launch(ctx);
}

// ---- the same machinery on the cooperative kernels (widths without a register instantiation: T1) ------------------------
// Tables are plain SoA ciphertext batches [2][fl][ne], ne = (2^W - 1) * nwin * nb, entry (d, w, base) at element
// (d-1)*nwin*nb + w*nb + base (coop_ops.cuh cq_tabprod), built with the batched squaring / composition launches:
//   B_w = ct^(2^(W w)) by W squarings per window; T[d] = T[d-1] o T[1] over all (w, base) at once for d = 2 .. 2^W - 1.
// dst[r][e] (row stride dst_n) = src[r][src_off + e] (row stride src_n) for e < cnt: a range of elements as its own batch
__global__ void __launch_bounds__(kBlock) k_cols_copy(int rows, size_t cnt, u32 *dst, size_t dst_n, const u32 *src, size_t src_n, size_t src_off) {
  size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (size_t)rows * cnt) return;
  size_t r = t / cnt, e = t % cnt;
  dst[r * dst_n + e] = src[r * src_n + src_off + e];
}
bool use_coop_tab(const cofhe_b200_ctx *ctx) { return ctx->coopn && use_coop(ctx) && (!ctx->fastn || ctx->opt.coop == 1); }
size_t coop_tab_entries(const MMShape &s) { return (size_t)s.n * s.m * s.nwin * ((1u << s.W) - 1); }
// scratch words of coop_tables
size_t coop_build_words(const cofhe_b200_ctx *ctx, const MMShape &s) {
  const size_t rows = 2 * (2 * (size_t)ctx->na + ctx->nc), nb = (size_t)s.n * s.m;
  return rows * (3 * nb * s.nwin + nb);
}
int cols_copy(cofhe_b200_ctx *ctx, int rows, size_t cnt, u32 *dst, size_t dst_n, const u32 *src, size_t src_n, cudaStream_t st) {
  const size_t tot = (size_t)rows * cnt;
  k_cols_copy<<<(unsigned)((tot + kBlock - 1) / kBlock), kBlock, 0, st>>>(rows, cnt, dst, dst_n, src, src_n, 0);
  return check_launch(ctx);
}
int coop_tables(cofhe_b200_ctx *ctx, const MMShape &s, const u32 *ct, u32 *tab, u32 *scratch, cudaStream_t st) {
  const int rows = (int)(2 * (2 * ctx->na + ctx->nc));
  const size_t nb = (size_t)s.n * s.m, nbw = nb * s.nwin, ne = coop_tab_entries(s);
  u32 *B = scratch, *X = B + (size_t)rows * nbw, *Y = X + (size_t)rows * nbw, *cur = Y + (size_t)rows * nbw;
  int rc = cols_copy(ctx, rows, nb, cur, nb, ct, nb, st);
  for (int w = 0; w < s.nwin && rc == COFHE_B200_OK; w++) {
    rc = cols_copy(ctx, rows, nb, B + (size_t)w * nb, nbw, cur, nb, st);
    if (w + 1 < s.nwin)
      for (int t = 0; t < s.W && rc == COFHE_B200_OK; t++) rc = launch_op(ctx, 1, 2, nb, cur, cur, cur, 0, nullptr, 0, 0, nullptr, 0, st);
  }
  if (rc != COFHE_B200_OK) return rc;
  rc = cols_copy(ctx, rows, nbw, tab, ne, B, nbw, st);
  const u32 *prev = B;
  const int nt = (1 << s.W) - 1;
  for (int d = 2; d <= nt && rc == COFHE_B200_OK; d++) {
    u32 *dst = (d & 1) ? X : Y;
    rc = launch_op(ctx, d == 2 ? 1 : 0, 2, nbw, prev, B, dst, 0, nullptr, 0, 0, nullptr, 0, st);
    if (rc == COFHE_B200_OK) rc = cols_copy(ctx, rows, nbw, tab + (size_t)(d - 1) * nbw, ne, dst, nbw, st);
    prev = dst;
  }
  return rc;
}
// out[i,k] = init o prod_j prod_w T[...]: one cooperative launch over (chunk, output), the chunks combined by batched
// compositions, flagged outputs redone from ct on the robust path.  part: chunks * 2 * fl * n*p words when chunks > 1.
int coop_mm_eval(cofhe_b200_ctx *ctx, const MMShape &s, const u32 *ct, const u32 *scal, const u32 *init, u32 *out, const u32 *tab, size_t ne,
                 u32 *part, cudaStream_t st) {
  Shape sh{(int)ctx->na, (int)ctx->nc};
  const size_t no = (size_t)s.n * s.p, fl = 2 * (size_t)ctx->na + ctx->nc;
  const size_t need = 2 * no * s.chunks + 2;
  if (need > ctx->tpbail_cap) {
    if (ctx->d_tpbail) cudaFree(ctx->d_tpbail);
    ctx->d_tpbail = nullptr; ctx->tpbail_cap = 0;
    if (cudaMalloc(&ctx->d_tpbail, need * sizeof(u32)) != cudaSuccess) { cudaGetLastError(); return COFHE_B200_ENOMEM; }
    ctx->tpbail_cap = need;
  }
  if (cudaMemsetAsync(ctx->d_tpbail, 0, 2 * sizeof(u32), st) != cudaSuccess) { cudaGetLastError(); return COFHE_B200_ECUDA; }
  const GSel gs = gsel(ctx, 2);
  u32 *dst = s.chunks > 1 ? part : out;
  // 256-thread blocks (8 warps aligned by the barriers, two blocks per SM) once every SM gets one: measured at T1, 8 192
  // fixed-base pairs 108 -> 68 ms, 2 048 pairs 39 -> 46 ms
  const size_t tp_groups = 2 * no * s.chunks, tp_lanes = ctx->coopn >= 53 ? 8 : (ctx->coopn >= 33 ? 4 : 2);
  const int tp_block = ctx->opt.coop_pb == 256 ? 256 : (ctx->opt.coop_pb == 128 ? 128 :
                       (tp_groups >= (size_t)(ctx->sms > 0 ? ctx->sms : 148) * (256 / tp_lanes) ? 256 : 128));
  COOP_DISPATCH(ctx->coopn, {
    typedef CoopKernels<CT, CL, CN, CCAP> K;
    K::tabprod(2, st, gs.G, gs.dbits, (int)ctx->nc, s, scal, tab, ne, init, dst, ctx->d_tpbail + 2, (int *)ctx->d_tpbail, tp_block);
  })
  int rc = check_launch(ctx);
  for (unsigned c = 1; c < s.chunks && rc == COFHE_B200_OK; c++)
    rc = launch_op(ctx, 0, 2, no, c == 1 ? part : out, part + (size_t)c * 2 * fl * no, out, 0, nullptr, 0, 0, nullptr, 0, st);
  if (rc != COFHE_B200_OK) return rc;
  CAP_DISPATCH(ctx->cap, k_mm_robust<CAP><<<32, 128, 0, st>>>((const Group<CAP> *)gs.G, sh, s, ct, scal, init, out, ctx->d_tpbail + 2,
                                                              (const int *)ctx->d_tpbail, (int *)ctx->d_tpbail + 1, ctx->d_err, gs.alt))
  return check_launch(ctx);
}

// (h^e, pk^e) for n exponents into a ciphertext-shaped batch `out`; part: scratch of mm_part_words(fb with p = n)
int fb_pow(cofhe_b200_ctx *ctx, size_t n, const u32 *exps, u32 el, u32 *out, u32 *part, cudaStream_t st) {
  MMShape s = ctx->fb;
  if ((int)el > s.el) return COFHE_B200_EINVAL;  // exponent wider than the tables were built for
  UseK in_cl_g(ctx);  // compact variant: h and pk are elements of Cl(DeltaK) (CL_HSM2k.inl:76-80, Cl_G :255)
  s.p = (unsigned)n;
  s.el = (int)el;
  if (ctx->fb_coop) return coop_mm_eval(ctx, s, ctx->d_fb_bases, exps, nullptr, out, ctx->d_fb_tab, ctx->fb_ne, nullptr, st);
  if (!ctx->fastn && use_coop(ctx)) {
    // no tables (set_bases ran while the cooperative path was off): (h^e, pk^e) as two exponentiations per element
    const int rows = (int)(2 * (2 * ctx->na + ctx->nc));
    k_bcast<<<grid_for(n), kBlock, 0, st>>>(n, rows, ctx->d_fb_bases, out);
    int rc = check_launch(ctx);
    if (rc != COFHE_B200_OK) return rc;
    return launch_op(ctx, 2, 2, n, out, out, out, 0, exps, (int)el, 1, nullptr, 0, st);
  }
  if (!ctx->fastn) return mm_robust_all(ctx, s, ctx->d_fb_bases, exps, nullptr, out, st);
  return mm_eval(ctx, s, ctx->d_fb_bases, exps, nullptr, out, ctx->d_fb_tab, part, st);
}

// forms of Cl(DeltaK) -> Cl(Delta), batched, in place allowed (compact variant only; the caller holds the context)
int lift_forms(cofhe_b200_ctx *ctx, size_t n, const u32 *in, u32 *out, cudaStream_t st) {
  if (!ctx->compact || !ctx->d_scheme) return COFHE_B200_ESTATE;
  if (n > ctx->lift_bytes) {
    if (ctx->d_lift) cudaFree(ctx->d_lift);
    ctx->d_lift = nullptr; ctx->lift_bytes = 0;
    if (cudaMalloc(&ctx->d_lift, n) != cudaSuccess) { cudaGetLastError(); return COFHE_B200_ENOMEM; }
    ctx->lift_bytes = n;
  }
  Shape s{(int)ctx->na, (int)ctx->nc};
  CAP_DISPATCH(ctx->cap, k_lift_2exp<CAP><<<grid_for(n), kBlock, 0, st>>>(s, n, ctx->k, in, out, ctx->d_lift, ctx->d_err))
  int rc = check_launch(ctx);
  if (rc != COFHE_B200_OK) return rc;
  // raise_to_power_M (CL_HSM2k.inl:641-648: k squarings in Cl(Delta)) as one exponentiation by 2^k
  u32 e2k[64];
  const int el = ctx->k / 32 + 1;
  for (int i = 0; i < el; i++) e2k[i] = 0;
  e2k[ctx->k / 32] = 1u << (ctx->k & 31);
  if (cudaMemcpyAsync(ctx->d_exp + 64, e2k, el * sizeof(u32), cudaMemcpyHostToDevice, st) != cudaSuccess) { cudaGetLastError(); return COFHE_B200_ECUDA; }
  const int force = ctx->gforce;
  ctx->gforce = 0;  // this step runs in Cl(Delta) whatever the caller is in the middle of
  rc = launch_op(ctx, 2, 1, n, out, out, out, 0, ctx->d_exp + 64, el, 0, nullptr, 0, st);
  ctx->gforce = force;
  if (rc != COFHE_B200_OK) return rc;
  CAP_DISPATCH(ctx->cap, k_lift_fix<CAP><<<grid_for(n), kBlock, 0, st>>>((const Group<CAP> *)ctx->d_group, (const Scheme<CAP> *)ctx->d_scheme,
                                                                      s, n, out, ctx->d_lift, ctx->d_err))
  return check_launch(ctx);
}

struct DevGuard {
  int prev;
  bool ok;
  explicit DevGuard(int dev) : prev(-1), ok(false) {
    if (cudaGetDevice(&prev) != cudaSuccess) return;
    ok = (prev == dev) || cudaSetDevice(dev) == cudaSuccess;
  }
  ~DevGuard() { if (ok && prev >= 0) cudaSetDevice(prev); }
};

// Entry guard of every call that issues work on `st`: the context lock, the device, and stream ordering.
struct Enter {
  std::unique_lock<std::recursive_mutex> lk;
  DevGuard g;
  bool ok;
  Enter(cofhe_b200_ctx *ctx, cudaStream_t st) : lk(ctx->mu), g(ctx->device), ok(g.ok) {
    if (!ok) return;
    if (ctx->busy && ctx->last != st) {
      if (cudaEventRecord(ctx->order, ctx->last) != cudaSuccess || cudaStreamWaitEvent(st, ctx->order, 0) != cudaSuccess) {
        cudaGetLastError();
        ok = false;
      }
    }
    ctx->last = st;
    ctx->busy = true;
  }
};
// Entry guard of the blocking _host calls (internal streams): waits for whatever the context still has in flight.
struct EnterHost {
  std::unique_lock<std::recursive_mutex> lk;
  DevGuard g;
  bool ok;
  explicit EnterHost(cofhe_b200_ctx *ctx) : lk(ctx->mu), g(ctx->device), ok(g.ok) {
    if (ok && ctx->busy) { ok = cudaStreamSynchronize(ctx->last) == cudaSuccess; ctx->busy = false; }
  }
};

}  // namespace

// ---- fixed scratch of cofhe_b200_add_ciphertexts_static ---------------------------------------------------------
// The one entry point that must not allocate keeps its scratch (the two counters and the list of declined elements that
// cofhe_b200_add_ciphertexts keeps in cudaMalloc'ed, grown-on-demand buffers) in this array.  Every GPU that loads the
// library holds its own copy (2 MiB + 256 bytes) until the process exits.  words [0, 64): counters (only [0] and [1]
// are used; the rest pads the list to a 256-byte boundary), words [64, 64 + 2 * COFHE_B200_STATIC_MAX_N): the list.
static __device__ __align__(256) u32 g_static_scratch[64 + 2 * (size_t)COFHE_B200_STATIC_MAX_N];

namespace {
// In host code `g_static_scratch` is only a handle for the runtime, not a device address: the address on a given device
// comes from cudaGetSymbolAddress with that device current.  Looked up once per device (cofhe_b200_create does it at
// set-up, the entry point finds it cached) -- the lookup allocates nothing.
constexpr int kStaticMaxDevices = 64;
std::mutex g_static_mu;
u32 *g_static_ptr[kStaticMaxDevices];
bool g_static_tried[kStaticMaxDevices];
// the caller has made `device` current
int static_scratch(int device, u32 **ptr) {
  if (device < 0 || device >= kStaticMaxDevices) return COFHE_B200_EINVAL;
  std::lock_guard<std::mutex> lk(g_static_mu);
  if (!g_static_tried[device]) {
    void *p = nullptr;
    cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
    if (e != cudaSuccess) { cudaGetLastError(); p = nullptr; }
    else g_static_tried[device] = true;  // a failure is retried by the next call
    g_static_ptr[device] = (u32 *)p;
  }
  *ptr = g_static_ptr[device];
  return *ptr ? COFHE_B200_OK : COFHE_B200_ECUDA;
}
}  // namespace

extern "C" {

const char *cofhe_b200_version(void) { return "cofhe_b200 0.1 sm_100a"; }

int cofhe_b200_create(cofhe_b200_ctx **out, int device, const uint32_t *abs_delta, uint32_t nd) {
  if (!out || !abs_delta || nd == 0) return COFHE_B200_EINVAL;
  *out = nullptr;
  u32 n = (u32)bn_norm(abs_delta, (int)nd);
  if (n == 0) return COFHE_B200_EINVAL;
  int cap = cap_for(n);
  if (!cap) return COFHE_B200_EINVAL;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) return COFHE_B200_ECUDA;
  DevGuard g(device);
  if (!g.ok) return COFHE_B200_ECUDA;
  if (getenv("COFHE_B200_STACK")) cudaDeviceSetLimit(cudaLimitStackSize, (size_t)atoi(getenv("COFHE_B200_STACK")));
  cofhe_b200_ctx *ctx = new (std::nothrow) cofhe_b200_ctx();
  if (!ctx) return COFHE_B200_ENOMEM;
  ctx->last = nullptr; ctx->busy = false; ctx->order = nullptr;
  ctx->opt.no_fast = getenv("COFHE_B200_NO_FAST") ? 1 : 0;
  ctx->opt.mm_w = env_opt("COFHE_B200_MM_W", 2, 8);
  ctx->opt.mm_chunks = env_opt("COFHE_B200_MM_CHUNKS", 1, 1 << 20);
  ctx->opt.mm_block = env_opt("COFHE_B200_MM_BLOCK", 128, 512);
  ctx->opt.fb_w = env_opt("COFHE_B200_FB_W", 2, 10);
  ctx->opt.host_staged = getenv("COFHE_B200_HOST_STAGED") ? 1 : 0;
  ctx->opt.host_loadc = getenv("COFHE_B200_HOST_LOADC") ? 1 : 0;
  ctx->opt.host_chunks = env_opt("COFHE_B200_HOST_CHUNKS", 0, 64);
  ctx->opt.host_trace = getenv("COFHE_B200_HOST_TRACE") ? 1 : 0;
  ctx->opt.coop = env_opt("COFHE_B200_COOP", 0, 1);
  ctx->opt.coop_mb = env_opt("COFHE_B200_COOP_MB", 3, 24);
  ctx->opt.coop_pb = env_opt("COFHE_B200_COOP_PB", 4, 256);
  ctx->opt.fast_block = env_opt("COFHE_B200_FAST_BLOCK", 128, 512);
  ctx->opt.fast_flat = env_opt("COFHE_B200_FAST_FLAT", 0, 1);
  ctx->device = device;
  ctx->sms = 0;
  if (cudaDeviceGetAttribute(&ctx->sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) { cudaGetLastError(); ctx->sms = 0; }
  ctx->cap = cap;
  ctx->nd = n;
  int dbits = bn_bits(abs_delta, (int)n);
  ctx->na = (u32)((dbits / 2 + 1 + 31) / 32);
  ctx->nc = n;
  ctx->launches = 0;
  ctx->dbits = (u32)dbits;
  for (u32 i = 0; i < n; i++) ctx->h_absd[i] = abs_delta[i];
  ctx->fastn = fast_n_for(ctx->na, cap, ctx->opt.no_fast != 0);
  ctx->coopn = (COOP_HAVE(ctx->na) && coop_cap((int)ctx->na) == cap && !getenv("COFHE_B200_NO_COOP")) ? (int)ctx->na : 0;
  ctx->d_bail = nullptr; ctx->d_nbail = nullptr; ctx->bail_cap = 0; ctx->flip = 0;
  memset(ctx->lane, 0, sizeof(ctx->lane));
  ctx->d_stage = nullptr; ctx->stage_bytes = 0;
  ctx->d_work = nullptr; ctx->work_bytes = 0;
  ctx->d_tmp = nullptr; ctx->tmp_bytes = 0;
  ctx->d_scheme = nullptr; ctx->k = 0; ctx->kl = 0; ctx->d_exp = nullptr;
  ctx->d_fb_bases = nullptr; ctx->d_fb_tab = nullptr; ctx->fb_set = 0;
  ctx->d_group = nullptr; ctx->d_groupK = nullptr; ctx->compact = 0; ctx->dbitsK = 0; ctx->gforce = 0;
  ctx->d_err = nullptr;
  ctx->xfer = nullptr;
  int rc = COFHE_B200_OK;
  CAP_DISPATCH(cap, {
    Group<CAP> *G = new (std::nothrow) Group<CAP>();
    if (!G) { delete ctx; return COFHE_B200_ENOMEM; }
    memset(G, 0, sizeof(*G));
    G->absD.n = (int)n; G->absD.neg = 0;
    for (u32 i = 0; i < n; i++) G->absD.d[i] = abs_delta[i];
    G->dbits = dbits;
    // two groups back to back: Delta, then DeltaK (a copy of the first until cofhe_b200_set_compact)
    if (cudaMalloc(&ctx->d_group, 2 * sizeof(*G)) != cudaSuccess ||
        cudaMemcpy(ctx->d_group, G, sizeof(*G), cudaMemcpyHostToDevice) != cudaSuccess ||
        cudaMemcpy((char *)ctx->d_group + sizeof(*G), G, sizeof(*G), cudaMemcpyHostToDevice) != cudaSuccess)
      rc = COFHE_B200_ECUDA;
    ctx->d_groupK = ctx->d_group ? (void *)((char *)ctx->d_group + sizeof(*G)) : nullptr;
    delete G;
  })
  if (rc == COFHE_B200_OK && (cudaMalloc(&ctx->d_err, sizeof(int)) != cudaSuccess ||
                              cudaMemset(ctx->d_err, 0, sizeof(int)) != cudaSuccess ||
                              cudaMalloc(&ctx->d_nbail, 2 * sizeof(int)) != cudaSuccess ||
                              cudaMalloc(&ctx->d_exp, 128 * sizeof(u32)) != cudaSuccess ||
                              cudaMemset(ctx->d_nbail, 0, 2 * sizeof(int)) != cudaSuccess ||
                              cudaEventCreateWithFlags(&ctx->order, cudaEventDisableTiming) != cudaSuccess ||
                              cudaStreamCreateWithFlags(&ctx->xfer, cudaStreamNonBlocking) != cudaSuccess))
    rc = COFHE_B200_ECUDA;
  if (rc != COFHE_B200_OK) { cofhe_b200_destroy(ctx); return rc; }
  { u32 *unused = nullptr; (void)static_scratch(device, &unused); }  // address of the fixed scratch on this device (no allocation)
  *out = ctx;
  return COFHE_B200_OK;
}

void cofhe_b200_destroy(cofhe_b200_ctx *ctx) {
  if (!ctx) return;
  DevGuard g(ctx->device);
  if (ctx->d_group) cudaFree(ctx->d_group);
  if (ctx->d_err) cudaFree(ctx->d_err);
  if (ctx->d_bail) cudaFree(ctx->d_bail);
  if (ctx->d_nbail) cudaFree(ctx->d_nbail);
  if (ctx->d_stage) cudaFree(ctx->d_stage);
  if (ctx->d_work) cudaFree(ctx->d_work);
  if (ctx->d_tmp) cudaFree(ctx->d_tmp);
  if (ctx->d_powtab) cudaFree(ctx->d_powtab);
  if (ctx->d_lift) cudaFree(ctx->d_lift);
  if (ctx->d_tpbail) cudaFree(ctx->d_tpbail);
  if (ctx->d_scheme) cudaFree(ctx->d_scheme);
  if (ctx->d_exp) cudaFree(ctx->d_exp);
  if (ctx->d_fb_bases) cudaFree(ctx->d_fb_bases);
  if (ctx->d_fb_tab) cudaFree(ctx->d_fb_tab);
  for (auto &b : ctx->pool) cudaFree(b.p);
  if (ctx->xfer) cudaStreamDestroy(ctx->xfer);
  if (ctx->order) cudaEventDestroy(ctx->order);
  for (auto &L : ctx->lane) {
    if (L.d_bail) cudaFree(L.d_bail);
    if (L.d_nbail) cudaFree(L.d_nbail);
    if (L.st) cudaStreamDestroy(L.st);
  }
  delete ctx;
}

int cofhe_b200_shape(const cofhe_b200_ctx *ctx, uint32_t *na, uint32_t *nc, uint32_t *nd) {
  if (!ctx) return COFHE_B200_EINVAL;
  if (na) *na = ctx->na;
  if (nc) *nc = ctx->nc;
  if (nd) *nd = ctx->nd;
  return COFHE_B200_OK;
}

int cofhe_b200_status(cofhe_b200_ctx *ctx, void *stream) {
  if (!ctx) return COFHE_B200_EINVAL;
  cudaStream_t st = (cudaStream_t)stream;
  Enter g(ctx, st);
  if (!g.ok) return COFHE_B200_ECUDA;
  int h = 0;
  CUDA_TRY(cudaMemcpyAsync(&h, ctx->d_err, sizeof(int), cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  ctx->busy = false;  // everything the context issued is ordered before `st`, which is idle now
  if (h) {
    CUDA_TRY(cudaMemsetAsync(ctx->d_err, 0, sizeof(int), st));
    CUDA_TRY(cudaStreamSynchronize(st));
    return COFHE_B200_EARITH;
  }
  return COFHE_B200_OK;
}

uint64_t cofhe_b200_launch_count(const cofhe_b200_ctx *ctx) { return ctx ? ctx->launches : 0; }
int cofhe_b200_fast_limbs(const cofhe_b200_ctx *ctx) { return ctx ? (ctx->fastn ? ctx->fastn : ctx->coopn) : 0; }
int cofhe_b200_coop_limbs(const cofhe_b200_ctx *ctx) { return ctx ? ctx->coopn : 0; }
int cofhe_b200_reg_limbs(const cofhe_b200_ctx *ctx) { return ctx ? ctx->fastn : 0; }

int cofhe_b200_set_option(cofhe_b200_ctx *ctx, const char *name, long value) {
  if (!ctx || !name) return COFHE_B200_EINVAL;
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  struct { const char *n; long *p; } tab[] = {
      {"mm_w", &ctx->opt.mm_w}, {"mm_chunks", &ctx->opt.mm_chunks}, {"mm_block", &ctx->opt.mm_block}, {"fb_w", &ctx->opt.fb_w},
      {"host_staged", &ctx->opt.host_staged}, {"host_loadc", &ctx->opt.host_loadc}, {"host_chunks", &ctx->opt.host_chunks},
      {"host_trace", &ctx->opt.host_trace}, {"coop", &ctx->opt.coop}, {"coop_mb", &ctx->opt.coop_mb}, {"coop_pb", &ctx->opt.coop_pb}, {"fast_block", &ctx->opt.fast_block}, {"fast_flat", &ctx->opt.fast_flat}};
  for (auto &t : tab)
    if (!strcmp(t.n, name)) { *t.p = value; return COFHE_B200_OK; }
  return COFHE_B200_EINVAL;
}

int cofhe_b200_nucomp(cofhe_b200_ctx *ctx, size_t n, const uint32_t *f1, const uint32_t *f2, uint32_t *out, int negf2,
                      void *stream) {
  if (!ctx || (n && (!f1 || !f2 || !out))) return COFHE_B200_EINVAL;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  return launch_op(ctx, 0, 1, n, f1, f2, out, negf2, nullptr, 0, 0, nullptr, 0, (cudaStream_t)stream);
}

int cofhe_b200_nudupl(cofhe_b200_ctx *ctx, size_t n, const uint32_t *f, uint32_t *out, void *stream) {
  if (!ctx || (n && (!f || !out))) return COFHE_B200_EINVAL;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  return launch_op(ctx, 1, 1, n, f, f, out, 0, nullptr, 0, 0, nullptr, 0, (cudaStream_t)stream);
}

int cofhe_b200_reduce(cofhe_b200_ctx *ctx, size_t n, const uint32_t *f, uint32_t na_in, uint32_t nc_in, uint32_t *out,
                      void *stream) {
  if (!ctx || (n && (!f || !out)) || na_in == 0 || nc_in == 0) return COFHE_B200_EINVAL;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  Shape sin{(int)na_in, (int)nc_in}, sout{(int)ctx->na, (int)ctx->nc};
  CAP_DISPATCH(ctx->cap, k_reduce<CAP><<<grid_for(n), kBlock, 0, (cudaStream_t)stream>>>((const Group<CAP> *)ctx->d_group, sin,
                                                                                         sout, n, f, out, ctx->d_err))
  return check_launch(ctx);
}

int cofhe_b200_nupow(cofhe_b200_ctx *ctx, size_t n, const uint32_t *f, const uint32_t *exp, uint32_t el, int per_elem,
                     const uint8_t *expneg, int shared_neg, uint32_t *out, void *stream) {
  if (!ctx || (n && (!f || !out || !exp)) || el == 0) return COFHE_B200_EINVAL;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  return launch_op(ctx, 2, 1, n, f, f, out, 0, exp, (int)el, per_elem, expneg, shared_neg, (cudaStream_t)stream);
}

int cofhe_b200_add_ciphertexts(cofhe_b200_ctx *ctx, size_t n, const uint32_t *ct1, const uint32_t *ct2, uint32_t *out,
                               void *stream) {
  if (!ctx || (n && (!ct1 || !ct2 || !out))) return COFHE_B200_EINVAL;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  return launch_op(ctx, 0, 2, n, ct1, ct2, out, 0, nullptr, 0, 0, nullptr, 0, (cudaStream_t)stream);
}

// cofhe_b200_add_ciphertexts without any allocation: the same kernels, their scratch in g_static_scratch.
int cofhe_b200_add_ciphertexts_static(cofhe_b200_ctx *ctx, size_t n, const uint32_t *ct1, const uint32_t *ct2, uint32_t *out,
                                      void *stream) {
  if (!ctx || (n && (!ct1 || !ct2 || !out))) return COFHE_B200_EINVAL;
  if (n > (size_t)COFHE_B200_STATIC_MAX_N) return COFHE_B200_ETOOBIG;
  if (n == 0) return COFHE_B200_OK;
  cudaStream_t st = (cudaStream_t)stream;
  Enter g(ctx, st);
  if (!g.ok) return COFHE_B200_ECUDA;
  u32 *scr = nullptr;
  int rc = static_scratch(ctx->device, &scr);
  if (rc != COFHE_B200_OK) return rc;
  int *cnt = (int *)scr, *cnt_next = cnt + 1;
  u32 *bail = scr + 64;
  const unsigned ny = 2;
  Shape s{(int)ctx->na, (int)ctx->nc};
  dim3 grid(grid_for(n), ny);
  const bool coop = use_coop(ctx);
  const GSel gs = gsel(ctx, ny);
  if (!ctx->fastn && !coop) {  // the generic kernel keeps no list
    CAP_DISPATCH(ctx->cap, k_compose<CAP><<<grid, kBlock, 0, st>>>((const Group<CAP> *)gs.G, s, n, ct1, ct2, out, 0, ctx->d_err, gs.alt))
    return check_launch(ctx);
  }
  // the array keeps what the previous call left: the counters start every call at zero
  if (cudaMemsetAsync(scr, 0, 64 * sizeof(u32), st) != cudaSuccess) { cudaGetLastError(); return COFHE_B200_ECUDA; }
  int dbits = gs.dbits, nc = (int)ctx->nc;
  if (coop) {
    COOP_DISPATCH(ctx->coopn, {
      typedef CoopKernels<CT, CL, CN, CCAP> K;
      K::compose(false, ny, st, gs.G, dbits, nc, n, ct1, ct2, out, 0, bail, cnt, ctx->opt.coop_mb > 0 ? (int)ctx->opt.coop_mb : 14);
    })
  } else {
    FAST_DISPATCH(ctx->fastn, {
      typedef FastKernels<FN, FCAP> K;
      K::compose(false, grid, st, gs.G, dbits, nc, n, ct1, ct2, out, 0, bail, cnt);
    })
  }
  rc = check_launch(ctx);
  if (rc != COFHE_B200_OK) return rc;
  unsigned rb = (unsigned)((size_t)ny * n < 4096 ? ((size_t)ny * n + kBlock - 1) / kBlock : 32);
  CAP_DISPATCH(ctx->cap, k_redo<CAP><<<rb, kBlock, 0, st>>>((const Group<CAP> *)gs.G, s, n, 0, ct1, ct2, out, 0, nullptr, 0, 0,
                                                          nullptr, 0, bail, cnt, cnt_next, ctx->d_err, gs.alt))
  return check_launch(ctx);
}

int cofhe_b200_scal_matmul(cofhe_b200_ctx *ctx, size_t n, size_t m, size_t p, const uint32_t *ct, const uint32_t *scal,
                           uint32_t el, uint32_t ebits, const uint32_t *init, uint32_t *out, void *stream) {
  if (!ctx || el == 0 || !init || ((n && p) && !out) || ((n && m) && !ct) || ((m && p) && !scal)) return COFHE_B200_EINVAL;
  if (n == 0 || p == 0) return COFHE_B200_OK;
  if ((n * p) >> 30 || (n * m) >> 30 || (m * p) >> 30) return COFHE_B200_EINVAL;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  cudaStream_t st = (cudaStream_t)stream;
  MMShape s;
  s.n = (unsigned)n; s.m = (unsigned)m; s.p = (unsigned)p; s.el = (int)el;
  if (ebits == 0 || ebits > 32 * el) ebits = 32 * el;
  // Window width: every base serves p outputs, so a wider window (fewer compositions per output, a larger table per
  // base) pays off with p.  Cost per base in compositions: p * nwin * (1 - 2^-W) for the outputs, nwin * (2^W - 1) table
  // entries (weighted 2: measured on B200, W = 5 and 6 tie at p = 256 for 17-bit scalars, 6 wins for 128-bit ones)
  // and nwin * W squarings for the chain; the table must also fit comfortably in HBM.
  {
    size_t fm = 0, tm = 0;
    if (cudaMemGetInfo(&fm, &tm) != cudaSuccess) { cudaGetLastError(); fm = (size_t)8 << 30; }
    const bool ctab = use_coop_tab(ctx);
    const double te_bytes = ctab ? 4.0 * (2 * ctx->na + ctx->nc) : 4.0 * (((3 * ctx->fastn + 1) + 3) & ~3);
    double best = 0;
    s.W = 4;
    for (int W = 2; W <= 8; W++) {
      double nwin = (double)((ebits + W - 1) / W), ent = (double)((1u << W) - 1);
      double bytes = 2.0 * (double)n * (double)m * nwin * ent * te_bytes;
      if ((ctx->fastn || ctab) && bytes > 0.4 * (double)fm && W > 2) continue;
      double cost = (double)p * nwin * (1.0 - 1.0 / (double)(1u << W)) + 2.0 * nwin * ent + nwin * W;
      if (best == 0 || cost < best) { best = cost; s.W = W; }
    }
  }
  if (ctx->opt.mm_w >= 2) s.W = (int)ctx->opt.mm_w;
  s.nwin = (int)((ebits + s.W - 1) / s.W);
  s.chunks = 1; s.jper = (unsigned)m; s.block = 0;
  size_t no = n * p;
  if (use_coop_tab(ctx) && m > 0) {
    // cooperative kernels: SoA tables, one lane group per (chunk, output); enough chunks of the j range to fill the machine
    // (3 blocks of 128 threads per SM), each extra chunk costs one batched composition per output
    const size_t sms = (size_t)(ctx->sms > 0 ? ctx->sms : 148);
    const size_t lanes = ctx->coopn >= 53 ? 8 : (ctx->coopn >= 33 ? 4 : 2);
    const size_t slots = sms * 3 * (kCoopBlock / lanes);
    size_t ch = (slots + 2 * no - 1) / (2 * no);
    if (ch > m) ch = m;
    if (ch < 1) ch = 1;
    if (ctx->opt.mm_chunks >= 1 && (size_t)ctx->opt.mm_chunks <= m) ch = (size_t)ctx->opt.mm_chunks;
    s.jper = (unsigned)((m + ch - 1) / ch);
    s.chunks = (unsigned)((m + s.jper - 1) / s.jper);
    const size_t fl = 2 * (size_t)ctx->na + ctx->nc, ne = coop_tab_entries(s);
    const size_t tw = 2 * fl * ne, bw = coop_build_words(ctx, s), pw = s.chunks > 1 ? (size_t)s.chunks * 2 * fl * no : 0;
    int rc = ensure_work(ctx, (tw + (bw > pw ? bw : pw)) * sizeof(u32));
    if (rc != COFHE_B200_OK) return rc;
    u32 *tab = ctx->d_work, *rest = tab + tw;  // the build scratch is dead when the partial products are written
    rc = coop_tables(ctx, s, ct, tab, rest, st);
    if (rc != COFHE_B200_OK) return rc;
    return coop_mm_eval(ctx, s, ct, scal, init, out, tab, ne, rest, st);
  }
  if (!ctx->fastn || m == 0) return mm_robust_all(ctx, s, ct, scal, init, out, st);
  // split the j range so that small n*p still fills the machine
  // k_mm_accum: blocks of `block` threads, 512 / block of them resident per SM (128 registers), all equally long: pick
  // the chunk count whose grid is closest to a whole number of waves (rounds * columns-per-chunk = the time of the
  // launch); each extra chunk costs one more composition per output in k_mm_final
  size_t ch = 1;
  {
    const size_t sms = (size_t)(ctx->sms > 0 ? ctx->sms : 148);
    size_t chmax = m / 2 < 1 ? 1 : m / 2;
    if (chmax > 64) chmax = 64;
    s.block = pick_block(2 * no * chmax, ctx->sms);  // full blocks whenever chunking can fill the machine with them
    if (ctx->opt.mm_block == 128 || ctx->opt.mm_block == 256 || ctx->opt.mm_block == 512) s.block = (int)ctx->opt.mm_block;
    const size_t slots = sms * (size_t)(512 / s.block);
    const size_t bpc = 2 * ((no + s.block - 1) / s.block);  // blocks per chunk
    double best = 0;
    for (size_t c = 1; c <= chmax; c++) {
      size_t jper = (m + c - 1) / c, cc = (m + jper - 1) / jper;
      size_t rounds = (bpc * cc + slots - 1) / slots;
      double cost = (double)rounds * (double)jper * (double)(s.nwin) + (double)cc;
      if (c == 1 || cost < best) { best = cost; ch = c; }
    }
  }
  if (ctx->opt.mm_chunks >= 1 && (size_t)ctx->opt.mm_chunks <= m) ch = (size_t)ctx->opt.mm_chunks;
  s.jper = (unsigned)((m + ch - 1) / ch);
  s.chunks = (unsigned)((m + s.jper - 1) / s.jper);
  size_t tw = mm_tab_words(ctx, s), pw = mm_part_words(ctx, s);
  int rc = ensure_work(ctx, (tw + pw) * sizeof(u32));
  if (rc != COFHE_B200_OK) return rc;
  u32 *tab = ctx->d_work, *part = tab + tw;
  rc = mm_tables(ctx, s, ct, tab, st);
  if (rc != COFHE_B200_OK) return rc;
  return mm_eval(ctx, s, ct, scal, init, out, tab, part, st);
}

// ---- CL_HSM2k scheme level ---------------------------------------------------------------------------------------
int cofhe_b200_set_scheme(cofhe_b200_ctx *ctx, uint32_t k, const uint32_t *N, uint32_t nN) {
  if (!ctx || !N || nN == 0 || k == 0 || k > 256) return COFHE_B200_EINVAL;
  EnterHost g(ctx);
  if (!g.ok) return COFHE_B200_ECUDA;
  int n = bn_norm(N, (int)nN);
  if (n == 0) return COFHE_B200_EINVAL;
  // |Delta| must be 2^(2k+5) * N  (CL_HSM2k.inl:620-637: DeltaK = -8N, Delta = DeltaK * 2^(2(k+1)))
  {
    int nb = bn_bits(N, n);
    if ((u32)nb + 2 * k + 5 != ctx->dbits) return COFHE_B200_EINVAL;
    unsigned sh = 2 * k + 5;
    for (u32 i = 0; i < ctx->nd; i++) {
      // limb i of N << sh
      long lo = (long)i - (long)(sh >> 5);
      u32 v = 0;
      if (lo >= 0 && lo < n) v = N[lo] << (sh & 31);
      if ((sh & 31) && lo - 1 >= 0 && lo - 1 < n) v |= N[lo - 1] >> (32 - (sh & 31));
      if (v != ctx->h_absd[i]) return COFHE_B200_EINVAL;
    }
  }
  ctx->k = (int)k;
  ctx->kl = k <= 32 ? 1 : (k <= 128 ? 4 : 8);
  int rc = COFHE_B200_OK;
  CAP_DISPATCH(ctx->cap, {
    Scheme<CAP> *S = new (std::nothrow) Scheme<CAP>();
    if (!S) return COFHE_B200_ENOMEM;
    memset(S, 0, sizeof(*S));
    Err e; e.code = 0;
    S->k = (int)k;
    S->N.n = n; S->N.neg = 0;
    for (int i = 0; i < n; i++) S->N.d[i] = N[i];
    z_shl(e, S->absDK, S->N, 3);
    Z<CAP> t, one;
    z_set_u32(one, 1);
    z_add(e, t, S->absDK, one);
    S->large = z_bits(t) <= (int)(2 * k) ? 1 : 0;
    if (e.code) rc = COFHE_B200_EINVAL;
    if (rc == COFHE_B200_OK && !ctx->d_scheme && cudaMalloc(&ctx->d_scheme, sizeof(*S)) != cudaSuccess) rc = COFHE_B200_ENOMEM;
    if (rc == COFHE_B200_OK && cudaMemcpy(ctx->d_scheme, S, sizeof(*S), cudaMemcpyHostToDevice) != cudaSuccess) rc = COFHE_B200_ECUDA;
    delete S;
  })
  return rc;
}

int cofhe_b200_power_of_f(cofhe_b200_ctx *ctx, size_t n, const uint32_t *m, uint32_t ml, uint32_t *out, void *stream) {
  if (!ctx || (n && (!m || !out)) || ml == 0) return COFHE_B200_EINVAL;
  if (!ctx->d_scheme) return COFHE_B200_ESTATE;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  Shape s{(int)ctx->na, (int)ctx->nc};
  SCHEME_DISPATCH(ctx, k_power_of_f<KL, CAP><<<grid_for(n), kBlock, 0, (cudaStream_t)stream>>>(
                           (const Group<CAP> *)ctx->d_group, (const Scheme<CAP> *)ctx->d_scheme, s, n, m, (int)ml, out, ctx->d_err))
  return check_launch(ctx);
}

int cofhe_b200_dlog(cofhe_b200_ctx *ctx, size_t n, const uint32_t *f, uint32_t *m, uint32_t ml, void *stream) {
  if (!ctx || (n && (!m || !f)) || ml == 0) return COFHE_B200_EINVAL;
  if (!ctx->d_scheme) return COFHE_B200_ESTATE;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  Shape s{(int)ctx->na, (int)ctx->nc};
  SCHEME_DISPATCH(ctx, k_dlog<KL, CAP><<<grid_for(n), kBlock, 0, (cudaStream_t)stream>>>(
                           (const Group<CAP> *)ctx->d_group, (const Scheme<CAP> *)ctx->d_scheme, s, n, f, m, (int)ml, ctx->d_err))
  return check_launch(ctx);
}

int cofhe_b200_set_compact(cofhe_b200_ctx *ctx, int on) {
  if (!ctx) return COFHE_B200_EINVAL;
  if (!ctx->d_scheme) return COFHE_B200_ESTATE;  // DeltaK = -8N comes from set_scheme
  EnterHost g(ctx);
  if (!g.ok) return COFHE_B200_ECUDA;
  int rc = COFHE_B200_OK;
  CAP_DISPATCH(ctx->cap, {
    Scheme<CAP> *S = new (std::nothrow) Scheme<CAP>();
    Group<CAP> *G = new (std::nothrow) Group<CAP>();
    if (!S || !G) { delete S; delete G; return COFHE_B200_ENOMEM; }
    memset(G, 0, sizeof(*G));
    if (cudaMemcpy(S, ctx->d_scheme, sizeof(*S), cudaMemcpyDeviceToHost) != cudaSuccess) rc = COFHE_B200_ECUDA;
    if (rc == COFHE_B200_OK) {
      G->absD = S->absDK;
      G->dbits = z_bits(S->absDK);
      if (cudaMemcpy(ctx->d_groupK, G, sizeof(*G), cudaMemcpyHostToDevice) != cudaSuccess) rc = COFHE_B200_ECUDA;
      ctx->dbitsK = (u32)G->dbits;
    }
    delete S; delete G;
  })
  if (rc != COFHE_B200_OK) return rc;
  ctx->compact = on ? 1 : 0;
  ctx->fb_set = 0;  // h and pk change groups with the variant: the tables must be rebuilt (cofhe_b200_set_bases)
  return COFHE_B200_OK;
}

int cofhe_b200_lift(cofhe_b200_ctx *ctx, size_t n, const uint32_t *f, uint32_t *out, void *stream) {
  if (!ctx || (n && (!f || !out))) return COFHE_B200_EINVAL;
  if (!ctx->compact || !ctx->d_scheme) return COFHE_B200_ESTATE;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  return lift_forms(ctx, n, f, out, (cudaStream_t)stream);
}

int cofhe_b200_set_bases(cofhe_b200_ctx *ctx, const uint32_t *h, const uint32_t *pk, uint32_t ebits) {
  if (!ctx || !h || ebits == 0) return COFHE_B200_EINVAL;
  EnterHost g(ctx);
  if (!g.ok) return COFHE_B200_ECUDA;
  size_t fl = 2 * ctx->na + ctx->nc;
  if (!ctx->d_fb_bases && cudaMalloc(&ctx->d_fb_bases, 2 * fl * sizeof(u32)) != cudaSuccess) return COFHE_B200_ENOMEM;
  CUDA_TRY(cudaStreamSynchronize(ctx->xfer));
  CUDA_TRY(cudaMemcpy(ctx->d_fb_bases, h, fl * sizeof(u32), cudaMemcpyHostToDevice));
  CUDA_TRY(cudaMemcpy(ctx->d_fb_bases + fl, pk ? pk : h, fl * sizeof(u32), cudaMemcpyHostToDevice));
  MMShape s;
  s.n = 1; s.m = 1; s.p = 0; s.chunks = 1; s.jper = 1; s.block = 0;
  s.W = 8;
  if (ctx->opt.fb_w >= 2) s.W = (int)ctx->opt.fb_w;
  s.el = (int)((ebits + 31) / 32);
  s.nwin = (32 * s.el + s.W - 1) / s.W;
  ctx->fb = s;
  ctx->fb_set = 1;
  ctx->fb_coop = 0;
  if (use_coop_tab(ctx)) {
    // tables for the cooperative multi-exponentiation (the widths without register kernels: T1)
    const size_t ne = coop_tab_entries(s);
    if (ctx->d_fb_tab) { cudaFree(ctx->d_fb_tab); ctx->d_fb_tab = nullptr; }
    u32 *scratch = nullptr;
    if (cudaMalloc(&ctx->d_fb_tab, 2 * fl * ne * sizeof(u32)) != cudaSuccess ||
        cudaMalloc(&scratch, coop_build_words(ctx, s) * sizeof(u32)) != cudaSuccess) { cudaGetLastError(); return COFHE_B200_ENOMEM; }
    int rc;
    {
      UseK in_cl_g(ctx);
      rc = coop_tables(ctx, s, ctx->d_fb_bases, ctx->d_fb_tab, scratch, ctx->xfer);
    }
    if (rc == COFHE_B200_OK) rc = cofhe_b200_status(ctx, ctx->xfer);
    cudaFree(scratch);
    if (rc != COFHE_B200_OK) return rc;
    ctx->fb_coop = 1;
    ctx->fb_ne = ne;
    return COFHE_B200_OK;
  }
  if (!ctx->fastn) return COFHE_B200_OK;
  if (ctx->d_fb_tab) { cudaFree(ctx->d_fb_tab); ctx->d_fb_tab = nullptr; }
  if (cudaMalloc(&ctx->d_fb_tab, mm_tab_words(ctx, s) * sizeof(u32)) != cudaSuccess) { cudaGetLastError(); return COFHE_B200_ENOMEM; }
  int rc;
  {
    UseK in_cl_g(ctx);  // compact variant: h and pk are elements of Cl(DeltaK)
    rc = mm_tables(ctx, s, ctx->d_fb_bases, ctx->d_fb_tab, ctx->xfer);
  }
  if (rc != COFHE_B200_OK) return rc;
  return cofhe_b200_status(ctx, ctx->xfer);
}

int cofhe_b200_fixed_base_pow(cofhe_b200_ctx *ctx, size_t n, const uint32_t *exps, uint32_t el, uint32_t *out, void *stream) {
  if (!ctx || (n && (!exps || !out)) || el == 0 || n >> 30) return COFHE_B200_EINVAL;
  if (!ctx->fb_set) return COFHE_B200_ESTATE;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  MMShape s = ctx->fb;
  s.p = (unsigned)n;
  int rc = ensure_tmp(ctx, (ctx->fastn ? mm_part_words(ctx, s) : 1) * sizeof(u32));
  if (rc != COFHE_B200_OK) return rc;
  return fb_pow(ctx, n, exps, el, out, ctx->d_tmp, (cudaStream_t)stream);
}

// out[.][e] = (sign[e] ? -mag[e] : mag[e]) mod 2^k as ml_out limb rows: what Mpz::mod(m, pt, 2^k) gives on the host
__global__ void __launch_bounds__(kBlock) k_pt_mod2k(size_t n, int k, const u32 *mag, int ml, const unsigned char *sign, u32 *out, int ml_out) {
  size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  const bool neg = sign && sign[e];
  u32 carry = 1;
  for (int i = 0; i < ml_out; i++) {
    u32 w = i < ml ? mag[(size_t)i * n + e] : 0u;
    if (neg) { u32 v = ~w + carry; carry = (carry && v == 0) ? 1u : 0u; w = v; }
    int hi = k - 32 * i;  // bits of this limb below 2^k
    if (hi <= 0) w = 0;
    else if (hi < 32) w &= (1u << hi) - 1u;
    out[(size_t)i * n + e] = w;
  }
}
int cofhe_b200_plaintexts_mod2k(cofhe_b200_ctx *ctx, size_t n, const uint32_t *mag, uint32_t ml, const uint8_t *sign, uint32_t *out,
                                uint32_t ml_out, void *stream) {
  if (!ctx || (n && (!mag || !out)) || ml == 0 || ml_out == 0) return COFHE_B200_EINVAL;
  if (!ctx->d_scheme || ctx->k <= 0) return COFHE_B200_ESTATE;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  k_pt_mod2k<<<grid_for(n), kBlock, 0, (cudaStream_t)stream>>>(n, ctx->k, mag, (int)ml, sign, out, (int)ml_out);
  return check_launch(ctx);
}

// ---- float codec (codec.cuh): make_plaintext / get_float_from_plaintext of the reference, one thread per element -----
__global__ void __launch_bounds__(kBlock) k_codec_make(size_t n, int k, const float *x, u32 *mag, int ml, unsigned char *sign, int *err) {
  size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  u32 m[16];
  float v = x[e];
  if (!(v - v == 0.0f)) { atomicMax(err, 1); v = 0.0f; }  // NaN / infinity: the reference's mpf_set_d is undefined there
  int nl = ml < 16 ? ml : 16;
  int sg = codec_make_plaintext(v, k, m, 16);
  for (int i = nl; i < 16; i++) if (m[i]) atomicMax(err, 1);  // wider than the caller's limbs
  for (int i = 0; i < ml; i++) mag[(size_t)i * n + e] = i < 16 ? m[i] : 0u;
  if (sign) sign[e] = (unsigned char)sg;
  else if (sg) atomicMax(err, 1);
}
__global__ void __launch_bounds__(kBlock) k_codec_get(size_t n, int k, const u32 *mag, int ml, const unsigned char *sign, float *out) {
  size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  u32 m[64];
  int nl = ml < 64 ? ml : 64;
  for (int i = 0; i < nl; i++) m[i] = mag[(size_t)i * n + e];
  out[e] = codec_get_float(m, nl, sign ? sign[e] : 0, k);
}
int cofhe_b200_make_plaintexts(cofhe_b200_ctx *ctx, size_t n, const float *x, uint32_t *mag, uint32_t ml, uint8_t *sign, void *stream) {
  if (!ctx || (n && (!x || !mag)) || ml == 0) return COFHE_B200_EINVAL;
  if (!ctx->d_scheme || ctx->k <= 0) return COFHE_B200_ESTATE;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  k_codec_make<<<grid_for(n), kBlock, 0, (cudaStream_t)stream>>>(n, ctx->k, x, mag, (int)ml, sign, ctx->d_err);
  return check_launch(ctx);
}
int cofhe_b200_plaintexts_to_floats(cofhe_b200_ctx *ctx, size_t n, const uint32_t *mag, uint32_t ml, const uint8_t *sign, float *out,
                                    void *stream) {
  if (!ctx || (n && (!mag || !out)) || ml == 0 || ml > 64) return COFHE_B200_EINVAL;
  if (!ctx->d_scheme || ctx->k <= 0) return COFHE_B200_ESTATE;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  k_codec_get<<<grid_for(n), kBlock, 0, (cudaStream_t)stream>>>(n, ctx->k, mag, (int)ml, sign, out);
  return check_launch(ctx);
}

int cofhe_b200_encrypt(cofhe_b200_ctx *ctx, size_t n, const uint32_t *m, uint32_t ml, const uint32_t *r, uint32_t el, int shared_r,
                       uint32_t *out, void *stream) {
  if (!ctx || (n && (!m || !out)) || !r || ml == 0 || el == 0 || n >> 30) return COFHE_B200_EINVAL;
  if (!ctx->fb_set || !ctx->d_scheme) return COFHE_B200_ESTATE;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  cudaStream_t st = (cudaStream_t)stream;
  size_t fl = 2 * ctx->na + ctx->nc;
  size_t nr = shared_r ? 1 : n;
  MMShape s = ctx->fb;
  s.p = (unsigned)nr;
  size_t pw = ctx->fastn ? mm_part_words(ctx, s) : 0;
  // tmp: partial products | f^m forms [fl][n] | the shared (h^r, pk^r) [2][fl][1]
  int rc = ensure_tmp(ctx, (pw + fl * n + 2 * fl) * sizeof(u32));
  if (rc != COFHE_B200_OK) return rc;
  u32 *part = ctx->d_tmp, *fm = part + pw, *one = fm + fl * n;
  if (shared_r) {
    rc = fb_pow(ctx, 1, r, el, one, part, st);
    if (rc != COFHE_B200_OK) return rc;
    if (ctx->compact) {  // pk^r from Cl(DeltaK) into Cl(Delta) (CL_HSM_utils.inl:178-179, cpu_cryptosystem_tensor_ops.inl:11-12)
      rc = lift_forms(ctx, 1, one + fl, one + fl, st);
      if (rc != COFHE_B200_OK) return rc;
    }
    k_bcast<<<grid_for(n), kBlock, 0, st>>>(n, (int)(2 * fl), one, out);
    rc = check_launch(ctx);
  } else {
    rc = fb_pow(ctx, n, r, el, out, part, st);
    if (rc == COFHE_B200_OK && ctx->compact) rc = lift_forms(ctx, n, out + fl * n, out + fl * n, st);
  }
  if (rc != COFHE_B200_OK) return rc;
  rc = cofhe_b200_power_of_f(ctx, n, m, ml, fm, stream);
  if (rc != COFHE_B200_OK) return rc;
  return launch_op(ctx, 0, 1, n, out + fl * n, fm, out + fl * n, 0, nullptr, 0, 0, nullptr, 0, st);
}

/* out[e] = (ct.c1[e] o h^r, ct.c2[e] o pk^r): the re-randomisation step of the single-ciphertext add / scal / negate
 * (reference thirdparty/bicycl/src/bicycl/CL_HSM_utils.inl:188-246: c1 = h^r o ..., c2 = pk^r o ...), batched. */
int cofhe_b200_rerandomize(cofhe_b200_ctx *ctx, size_t n, const uint32_t *ct, const uint32_t *r, uint32_t el, int shared_r, uint32_t *out,
                           void *stream) {
  if (!ctx || (n && (!ct || !out)) || !r || el == 0 || n >> 30) return COFHE_B200_EINVAL;
  if (!ctx->fb_set) return COFHE_B200_ESTATE;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  cudaStream_t st = (cudaStream_t)stream;
  size_t fl = 2 * ctx->na + ctx->nc;
  size_t nr = shared_r ? 1 : n;
  MMShape s = ctx->fb;
  s.p = (unsigned)nr;
  size_t pw = ctx->fastn ? mm_part_words(ctx, s) : 0;
  // tmp: partial products | (h^r, pk^r) batch [2][fl][n] | the shared pair [2][fl][1]
  int rc = ensure_tmp(ctx, (pw + 2 * fl * n + 2 * fl) * sizeof(u32));
  if (rc != COFHE_B200_OK) return rc;
  u32 *part = ctx->d_tmp, *hr = part + pw, *one = hr + 2 * fl * n;
  if (shared_r) {
    rc = fb_pow(ctx, 1, r, el, one, part, st);
    if (rc != COFHE_B200_OK) return rc;
    if (ctx->compact) {  // CL_HSM_utils.inl:204-205, 235-236
      rc = lift_forms(ctx, 1, one + fl, one + fl, st);
      if (rc != COFHE_B200_OK) return rc;
    }
    k_bcast<<<grid_for(n), kBlock, 0, st>>>(n, (int)(2 * fl), one, hr);
    rc = check_launch(ctx);
  } else {
    rc = fb_pow(ctx, n, r, el, hr, part, st);
    if (rc == COFHE_B200_OK && ctx->compact) rc = lift_forms(ctx, n, hr + fl * n, hr + fl * n, st);
  }
  if (rc != COFHE_B200_OK) return rc;
  return launch_op(ctx, 0, 2, n, ct, hr, out, 0, nullptr, 0, 0, nullptr, 0, st);
}

int cofhe_b200_scal_ciphertexts(cofhe_b200_ctx *ctx, size_t n, const uint32_t *ct, const uint32_t *exp, uint32_t el, int per_elem,
                                const uint8_t *expneg, int shared_neg, uint32_t *out, void *stream) {
  if (!ctx || (n && (!ct || !out)) || !exp || el == 0) return COFHE_B200_EINVAL;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  return launch_op(ctx, 2, 2, n, ct, ct, out, 0, exp, (int)el, per_elem, expneg, shared_neg, (cudaStream_t)stream);
}

// out = ct o x^(2^k - 1): ct + negate(x), both without randomness (the first half of a Beaver multiplication)
static int beaver_sub(cofhe_b200_ctx *ctx, size_t n, const u32 *ct, const u32 *x, u32 *out, cudaStream_t st) {
  int rc = launch_op(ctx, 2, 2, n, x, x, out, 0, ctx->d_exp, (ctx->k + 31) / 32, 0, nullptr, 0, st);
  if (rc != COFHE_B200_OK) return rc;
  return launch_op(ctx, 0, 2, n, ct, out, out, 0, nullptr, 0, 0, nullptr, 0, st);
}

int cofhe_b200_beaver_open(cofhe_b200_ctx *ctx, size_t n, const uint32_t *ct1, const uint32_t *ct2, const uint32_t *a,
                           const uint32_t *b, uint32_t *out1, uint32_t *out2, void *stream) {
  if (!ctx || (n && (!ct1 || !ct2 || !a || !b || !out1 || !out2))) return COFHE_B200_EINVAL;
  if (!ctx->d_scheme || ctx->k <= 0 || ctx->k > 32 * 64) return COFHE_B200_ESTATE;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  cudaStream_t st = (cudaStream_t)stream;
  // the exponent of negate_ciphertext_tensor: make_plaintext(-1) = 2^k - 1
  u32 ones[64];
  int el = (ctx->k + 31) / 32;
  for (int i = 0; i < el; i++) ones[i] = 0xffffffffu;
  if (ctx->k & 31) ones[el - 1] = (1u << (ctx->k & 31)) - 1;
  CUDA_TRY(cudaMemcpyAsync(ctx->d_exp, ones, el * sizeof(u32), cudaMemcpyHostToDevice, st));
  int rc = beaver_sub(ctx, n, ct1, a, out1, st);
  if (rc != COFHE_B200_OK) return rc;
  return beaver_sub(ctx, n, ct2, b, out2, st);
}

int cofhe_b200_beaver_close(cofhe_b200_ctx *ctx, size_t n, const uint32_t *pt1, const uint32_t *pt2, uint32_t ml, const uint32_t *a,
                            const uint32_t *b, const uint32_t *c, const uint32_t *r, uint32_t el, uint32_t *out, void *stream) {
  if (!ctx || (n && (!pt1 || !pt2 || !a || !b || !c || !out)) || !r || ml == 0 || ml > 16 || el == 0) return COFHE_B200_EINVAL;
  if (!ctx->fb_set || !ctx->d_scheme) return COFHE_B200_ESTATE;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  cudaStream_t st = (cudaStream_t)stream;
  size_t cw = (size_t)2 * (2 * ctx->na + ctx->nc) * n;
  // work: b^pt1 | a^pt2 | pt1 * pt2
  int rc = ensure_work(ctx, (2 * cw + (size_t)ml * n) * sizeof(u32));
  if (rc != COFHE_B200_OK) return rc;
  u32 *t1 = ctx->d_work, *t2 = t1 + cw, *prod = t2 + cw;
  rc = launch_op(ctx, 2, 2, n, b, b, t1, 0, pt1, (int)ml, 1, nullptr, 0, st);
  if (rc != COFHE_B200_OK) return rc;
  rc = launch_op(ctx, 2, 2, n, a, a, t2, 0, pt2, (int)ml, 1, nullptr, 0, st);
  if (rc != COFHE_B200_OK) return rc;
  k_mul_lo<<<grid_for(n), kBlock, 0, st>>>(n, (int)ml, pt1, pt2, prod);
  rc = check_launch(ctx);
  if (rc != COFHE_B200_OK) return rc;
  rc = cofhe_b200_encrypt(ctx, n, prod, ml, r, el, 1, out, stream);
  if (rc != COFHE_B200_OK) return rc;
  rc = launch_op(ctx, 0, 2, n, t1, t2, t1, 0, nullptr, 0, 0, nullptr, 0, st);
  if (rc != COFHE_B200_OK) return rc;
  rc = launch_op(ctx, 0, 2, n, t1, c, t1, 0, nullptr, 0, 0, nullptr, 0, st);
  if (rc != COFHE_B200_OK) return rc;
  return launch_op(ctx, 0, 2, n, t1, out, out, 0, nullptr, 0, 0, nullptr, 0, st);
}

// pt3[.][3 i + c] = (a_i, b_i, a_i * b_i mod 2^(32 ml)) for c = 0, 1, 2: the [n, 3] plaintext tensor of
// BeaversTripletGenerator::generate (reference include/smpc/beavers_triplet_generation.hpp:19-27), row-major
__global__ void __launch_bounds__(kBlock) k_triplet_rows(size_t n, int ml, const u32 *a, const u32 *b, u32 *pt3) {
  size_t e = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (e >= n) return;
  u32 r[16];
  for (int i = 0; i < ml; i++) r[i] = 0;
  for (int i = 0; i < ml; i++) {
    unsigned long long c = 0;
    u32 ai = a[(size_t)i * n + e];
    for (int j = 0; i + j < ml; j++) {
      c += (unsigned long long)ai * b[(size_t)j * n + e] + r[i + j];
      r[i + j] = (u32)c;
      c >>= 32;
    }
  }
  for (int i = 0; i < ml; i++) {
    pt3[(size_t)i * 3 * n + 3 * e] = a[(size_t)i * n + e];
    pt3[(size_t)i * 3 * n + 3 * e + 1] = b[(size_t)i * n + e];
    pt3[(size_t)i * 3 * n + 3 * e + 2] = r[i];
  }
}
/* Beaver triplets: out = encrypt_tensor([[a_i, b_i, a_i b_i]]) -- 3n ciphertexts in the reference's [n, 3] row-major order,
 * ONE randomness for the tensor (reference include/smpc/beavers_triplet_generation.hpp:17-37; the cache of
 * include/smpc/smpc_client.hpp:17,218 is filled by calls of this). */
int cofhe_b200_beaver_triplets(cofhe_b200_ctx *ctx, size_t n, const uint32_t *a, const uint32_t *b, uint32_t ml, const uint32_t *r, uint32_t el,
                               uint32_t *out, void *stream) {
  if (!ctx || (n && (!a || !b || !out)) || !r || ml == 0 || ml > 16 || el == 0 || (3 * n) >> 30) return COFHE_B200_EINVAL;
  if (!ctx->fb_set || !ctx->d_scheme) return COFHE_B200_ESTATE;
  if (n == 0) return COFHE_B200_OK;
  {
    Enter g(ctx, (cudaStream_t)stream);
    if (!g.ok) return COFHE_B200_ECUDA;
    int rc = ensure_work(ctx, (size_t)ml * 3 * n * sizeof(u32));
    if (rc != COFHE_B200_OK) return rc;
    k_triplet_rows<<<grid_for(n), kBlock, 0, (cudaStream_t)stream>>>(n, (int)ml, a, b, ctx->d_work);
    rc = check_launch(ctx);
    if (rc != COFHE_B200_OK) return rc;
  }
  return cofhe_b200_encrypt(ctx, 3 * n, ctx->d_work, ml, r, el, 1, out, stream);
}

// ---- ciphertext x ciphertext MATRIX product: the n*m*p expansion around the Beaver multiplication and the reduction of
// its n*m*p products (reference include/smpc/ciphertext_multiplications.hpp:51-110) -------------------------------------
// out1[(i*m + j)*p + k] = ct1[i*m + j], out2[(i*m + j)*p + k] = ct2[j*p + k]   (rows = 2 * form limbs; :58-70)
__global__ void __launch_bounds__(kBlock) k_ctmm_expand(int rows, size_t n, size_t m, size_t p, const u32 *ct1, const u32 *ct2, u32 *out1,
                                                        u32 *out2) {
  const size_t nmp = n * m * p;
  size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (size_t)rows * nmp) return;
  size_t r = t / nmp, e = t % nmp, ij = e / p, k = e % p, j = ij % m;
  out1[t] = ct1[r * (n * m) + ij];
  out2[t] = ct2[r * (m * p) + j * p + k];
}
// slab j of the products (the n*p elements (i, k) of inner index j) into a batch of `cnt` slabs starting at slab j0:
// dst[r][(j - j0)*np + i*p + k] = prod[r][(i*m + j)*p + k]
__global__ void __launch_bounds__(kBlock) k_ctmm_slabs(int rows, size_t n, size_t m, size_t p, size_t j0, size_t cnt, const u32 *prod, u32 *dst) {
  const size_t np = n * p, tot = cnt * np;
  size_t t = (size_t)blockIdx.x * blockDim.x + threadIdx.x;
  if (t >= (size_t)rows * tot) return;
  size_t r = t / tot, e = t % tot, j = j0 + e / np, ik = e % np, i = ik / p, k = ik % p;
  dst[t] = prod[r * (n * m * p) + (i * m + j) * p + k];
}
int cofhe_b200_ct_matmul_expand(cofhe_b200_ctx *ctx, size_t n, size_t m, size_t p, const uint32_t *ct1, const uint32_t *ct2, uint32_t *out1,
                                uint32_t *out2, void *stream) {
  if (!ctx || !ct1 || !ct2 || !out1 || !out2 || n == 0 || m == 0 || p == 0 || (n * m * p) >> 30) return COFHE_B200_EINVAL;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  int rows = (int)(2 * (2 * ctx->na + ctx->nc));
  size_t tot = (size_t)rows * n * m * p;
  k_ctmm_expand<<<(unsigned)((tot + kBlock - 1) / kBlock), kBlock, 0, (cudaStream_t)stream>>>(rows, n, m, p, ct1, ct2, out1, out2);
  return check_launch(ctx);
}

// out[i*p + k] = init o prod_j prod[(i*m + j)*p + k]: a balanced tree over j -- every level is ONE batched composition launch over
// half of the remaining slabs (log2 m launches of decreasing size instead of m launches of n*p elements)
int cofhe_b200_ct_matmul_reduce(cofhe_b200_ctx *ctx, size_t n, size_t m, size_t p, const uint32_t *prod, const uint32_t *init, uint32_t *out,
                                void *stream) {
  if (!ctx || !prod || !init || !out || n == 0 || m == 0 || p == 0 || (n * m * p) >> 30) return COFHE_B200_EINVAL;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  cudaStream_t st = (cudaStream_t)stream;
  const int rows = (int)(2 * (2 * ctx->na + ctx->nc));
  const size_t np = n * p, half = (m + 1) / 2;
  // work: A | B | C (each up to `half` slabs) | E (the slab left over at odd levels, accumulated) | Z (init broadcast)
  int rc = ensure_work(ctx, ((size_t)rows * (3 * half + 2) * np) * sizeof(u32));
  if (rc != COFHE_B200_OK) return rc;
  u32 *A = ctx->d_work, *B = A + (size_t)rows * half * np, *C = B + (size_t)rows * half * np, *E = C + (size_t)rows * half * np,
      *Z = E + (size_t)rows * np;
  auto blocks = [](size_t tot) { return (unsigned)((tot + kBlock - 1) / kBlock); };
  k_bcast<<<grid_for(np), kBlock, 0, st>>>(np, rows, init, Z);
  ctx->launches++;
  bool have_e = false;
  size_t cur = m;
  const u32 *src = nullptr;  // level input: nullptr = the products themselves
  size_t src_n = 0;
  while (cur > 1) {
    size_t h = cur / 2;
    if (!src) {
      k_ctmm_slabs<<<blocks((size_t)rows * h * np), kBlock, 0, st>>>(rows, n, m, p, 0, h, prod, A);
      k_ctmm_slabs<<<blocks((size_t)rows * h * np), kBlock, 0, st>>>(rows, n, m, p, h, h, prod, B);
    } else {
      k_cols_copy<<<blocks((size_t)rows * h * np), kBlock, 0, st>>>(rows, h * np, A, h * np, src, src_n, 0);
      k_cols_copy<<<blocks((size_t)rows * h * np), kBlock, 0, st>>>(rows, h * np, B, h * np, src, src_n, h * np);
    }
    ctx->launches += 2;
    if (cur & 1) {
      // the odd slab out joins the side accumulator
      u32 *tmp = have_e ? (u32 *)out : E;  // out is free until the end: use it as the landing buffer of the slab
      if (!src) k_ctmm_slabs<<<blocks((size_t)rows * np), kBlock, 0, st>>>(rows, n, m, p, 2 * h, 1, prod, tmp);
      else k_cols_copy<<<blocks((size_t)rows * np), kBlock, 0, st>>>(rows, np, tmp, np, src, src_n, 2 * h * np);
      ctx->launches++;
      if (have_e) {
        rc = launch_op(ctx, 0, 2, np, E, tmp, E, 0, nullptr, 0, 0, nullptr, 0, st);
        if (rc != COFHE_B200_OK) return rc;
      }
      have_e = true;
    }
    rc = launch_op(ctx, 0, 2, h * np, A, B, C, 0, nullptr, 0, 0, nullptr, 0, st);
    if (rc != COFHE_B200_OK) return rc;
    src = C; src_n = h * np;
    cur = h;
  }
  const u32 *last = src;
  if (!src) {  // m == 1: the products are the single slab
    k_ctmm_slabs<<<blocks((size_t)rows * np), kBlock, 0, st>>>(rows, n, m, p, 0, 1, prod, A);
    ctx->launches++;
    last = A;
  }
  rc = launch_op(ctx, 0, 2, np, Z, last, Z, 0, nullptr, 0, 0, nullptr, 0, st);
  if (rc != COFHE_B200_OK) return rc;
  if (have_e) return launch_op(ctx, 0, 2, np, Z, E, out, 0, nullptr, 0, 0, nullptr, 0, st);
  k_cols_copy<<<blocks((size_t)rows * np), kBlock, 0, st>>>(rows, np, out, np, Z, np, 0);
  return check_launch(ctx);
}

int cofhe_b200_part_decrypt(cofhe_b200_ctx *ctx, size_t n, const uint32_t *ct, const uint32_t *share, uint32_t el, int share_neg,
                            uint32_t *out, void *stream) {
  if (!ctx || (n && (!ct || !out)) || !share || el == 0 || el > 64) return COFHE_B200_EINVAL;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  cudaStream_t st = (cudaStream_t)stream;
  CUDA_TRY(cudaMemcpyAsync(ctx->d_exp, share, el * sizeof(u32), cudaMemcpyHostToDevice, st));
  if (!ctx->compact) return launch_op(ctx, 2, 1, n, ct, ct, out, 0, ctx->d_exp, (int)el, 0, nullptr, share_neg, st);
  // compact variant: c1^share in Cl(DeltaK), then into Cl(Delta) (cpu_cryptosystem_distributed.inl:249-251)
  int rc;
  {
    UseK in_cl_g(ctx);
    rc = launch_op(ctx, 2, 1, n, ct, ct, out, 0, ctx->d_exp, (int)el, 0, nullptr, share_neg, st);
  }
  if (rc != COFHE_B200_OK) return rc;
  return lift_forms(ctx, n, out, out, st);
}

int cofhe_b200_combine(cofhe_b200_ctx *ctx, size_t n, uint32_t t, const uint32_t *ct, const uint32_t *ds, uint32_t *m, uint32_t ml,
                       void *stream) {
  if (!ctx || (n && (!ct || !ds || !m)) || t == 0 || ml == 0) return COFHE_B200_EINVAL;
  if (!ctx->d_scheme) return COFHE_B200_ESTATE;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  cudaStream_t st = (cudaStream_t)stream;
  size_t fl = 2 * ctx->na + ctx->nc;
  int rc = ensure_tmp(ctx, 2 * fl * n * sizeof(u32));
  if (rc != COFHE_B200_OK) return rc;
  u32 *d = ctx->d_tmp, *r = d + fl * n;
  // d = ds_0 o ds_1^-1 o ... o ds_{t-1}^-1   (compute_d, lambda = (1,-1,-1,...): cpu_cryptosystem_distributed.inl:214-242)
  const u32 *cur = ds;
  for (uint32_t i = 1; i < t; i++) {
    rc = launch_op(ctx, 0, 1, n, cur, ds + (size_t)i * fl * n, d, 1, nullptr, 0, 0, nullptr, 0, st);
    if (rc != COFHE_B200_OK) return rc;
    cur = d;
  }
  // r = c2 o d^-1 ; m = dlog_F(r)   (finalDecrypt :256-270)
  rc = launch_op(ctx, 0, 1, n, ct + fl * n, cur, r, 1, nullptr, 0, 0, nullptr, 0, st);
  if (rc != COFHE_B200_OK) return rc;
  return cofhe_b200_dlog(ctx, n, r, m, ml, stream);
}

int cofhe_b200_decrypt(cofhe_b200_ctx *ctx, size_t n, const uint32_t *ct, const uint32_t *sk, uint32_t el, uint32_t *m, uint32_t ml,
                       void *stream) {
  if (!ctx || (n && (!ct || !m)) || !sk || el == 0 || el > 64 || ml == 0) return COFHE_B200_EINVAL;
  if (!ctx->d_scheme) return COFHE_B200_ESTATE;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  cudaStream_t st = (cudaStream_t)stream;
  size_t fl = 2 * ctx->na + ctx->nc;
  int rc = ensure_tmp(ctx, 2 * fl * n * sizeof(u32));
  if (rc != COFHE_B200_OK) return rc;
  u32 *x = ctx->d_tmp, *r = x + fl * n;
  CUDA_TRY(cudaMemcpyAsync(ctx->d_exp, sk, el * sizeof(u32), cudaMemcpyHostToDevice, st));
  {
    UseK in_cl_g(ctx);  // no effect outside the compact variant
    rc = launch_op(ctx, 2, 1, n, ct, ct, x, 0, ctx->d_exp, (int)el, 0, nullptr, 0, st);   // c1^sk
  }
  if (rc == COFHE_B200_OK && ctx->compact) rc = lift_forms(ctx, n, x, x, st);            // CL_HSM_utils.inl:132-133
  if (rc != COFHE_B200_OK) return rc;
  rc = launch_op(ctx, 0, 1, n, ct + fl * n, x, r, 1, nullptr, 0, 0, nullptr, 0, st);     // c2 o (c1^sk)^-1
  if (rc != COFHE_B200_OK) return rc;
  return cofhe_b200_dlog(ctx, n, r, m, ml, stream);
}

// ---- device memory helpers for callers that do not link the CUDA runtime (the header class) -----------------------
int cofhe_b200_dev_alloc(cofhe_b200_ctx *ctx, size_t bytes, void **ptr) {
  if (!ctx || !ptr) return COFHE_B200_EINVAL;
  *ptr = nullptr;
  if (bytes == 0) return COFHE_B200_OK;
  DevGuard g(ctx->device);
  if (!g.ok) return COFHE_B200_ECUDA;
  {
    // recycled block: freed buffers are kept (cudaMalloc / cudaFree synchronise the device and cost ~0.1-1 ms each, which
    // would dominate chained calls on device-resident tensors); work on a context is ordered (one stream at a time, stream
    // changes wait for the previous stream), so a block handed out again is never still in use by an earlier call
    std::lock_guard<std::recursive_mutex> lk(ctx->mu);
    size_t best = ctx->pool.size();
    for (size_t i = 0; i < ctx->pool.size(); i++)
      if (ctx->pool[i].bytes >= bytes && ctx->pool[i].bytes <= 2 * bytes + 4096 &&
          (best == ctx->pool.size() || ctx->pool[i].bytes < ctx->pool[best].bytes)) best = i;
    if (best != ctx->pool.size()) {
      *ptr = ctx->pool[best].p;
      ctx->live.push_back(ctx->pool[best]);
      ctx->pool_bytes -= ctx->pool[best].bytes;
      ctx->pool.erase(ctx->pool.begin() + best);
      return COFHE_B200_OK;
    }
  }
  if (cudaMalloc(ptr, bytes) != cudaSuccess) {
    cudaGetLastError();
    // out of memory: drop the recycled blocks and try once more
    {
      std::lock_guard<std::recursive_mutex> lk(ctx->mu);
      for (auto &b : ctx->pool) cudaFree(b.p);
      ctx->pool.clear(); ctx->pool_bytes = 0;
    }
    if (cudaMalloc(ptr, bytes) != cudaSuccess) { cudaGetLastError(); *ptr = nullptr; return COFHE_B200_ENOMEM; }
  }
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  ctx->live.push_back({*ptr, bytes});
  return COFHE_B200_OK;
}
void cofhe_b200_dev_free(cofhe_b200_ctx *ctx, void *ptr) {
  if (!ctx || !ptr) return;
  DevGuard g(ctx->device);
  std::lock_guard<std::recursive_mutex> lk(ctx->mu);
  for (size_t i = 0; i < ctx->live.size(); i++)
    if (ctx->live[i].p == ptr) {
      cofhe_b200_ctx::Block b = ctx->live[i];
      ctx->live.erase(ctx->live.begin() + i);
      if (ctx->pool_bytes + b.bytes <= ((size_t)4 << 30) && ctx->pool.size() < 64) { ctx->pool.push_back(b); ctx->pool_bytes += b.bytes; }
      else cudaFree(b.p);
      return;
    }
  cudaFree(ptr);
}
int cofhe_b200_host_alloc(size_t bytes, void **ptr) {
  if (!ptr) return COFHE_B200_EINVAL;
  *ptr = nullptr;
  if (bytes == 0) return COFHE_B200_OK;
  if (cudaMallocHost(ptr, bytes) != cudaSuccess) { cudaGetLastError(); *ptr = nullptr; return COFHE_B200_ENOMEM; }
  return COFHE_B200_OK;
}
void cofhe_b200_host_free(void *ptr) { if (ptr) cudaFreeHost(ptr); }
void *cofhe_b200_stream(cofhe_b200_ctx *ctx) { return ctx ? (void *)ctx->xfer : nullptr; }
int cofhe_b200_upload(cofhe_b200_ctx *ctx, void *dst, const void *src, size_t bytes, void *stream) {
  if (!ctx || (bytes && (!dst || !src))) return COFHE_B200_EINVAL;
  if (bytes == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyHostToDevice, (cudaStream_t)stream));
  return COFHE_B200_OK;
}
int cofhe_b200_download(cofhe_b200_ctx *ctx, void *dst, const void *src, size_t bytes, void *stream) {
  if (!ctx || (bytes && (!dst || !src))) return COFHE_B200_EINVAL;
  if (bytes == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToHost, (cudaStream_t)stream));
  return COFHE_B200_OK;
}
int cofhe_b200_copy(cofhe_b200_ctx *ctx, void *dst, const void *src, size_t bytes, void *stream) {
  if (!ctx || (bytes && (!dst || !src))) return COFHE_B200_EINVAL;
  if (bytes == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  CUDA_TRY(cudaMemcpyAsync(dst, src, bytes, cudaMemcpyDeviceToDevice, (cudaStream_t)stream));
  return COFHE_B200_OK;
}
int cofhe_b200_device_count(void) {
  int c = 0;
  if (cudaGetDeviceCount(&c) != cudaSuccess) { cudaGetLastError(); return 0; }
  return c;
}

/* ---- binary tensor wire format (reference cpu_cryptosystem.inl:322-520) <-> SoA batches, on the device -------- */
int cofhe_b200_wire_header(const uint8_t *wire, size_t bytes, uint32_t *ndim, uint32_t *dims, uint32_t max_dims, size_t *n) {
  if (!wire || !ndim || !n || bytes < 4) return COFHE_B200_EINVAL;
  uint32_t nd;
  memcpy(&nd, wire, 4);
  if (nd > max_dims || bytes < 4 + 4 * (size_t)nd) return COFHE_B200_EINVAL;
  size_t cnt = 1;
  for (uint32_t i = 0; i < nd; i++) {
    uint32_t d;
    memcpy(&d, wire + 4 + 4 * (size_t)i, 4);
    if (dims) dims[i] = d;
    if (d && cnt > ((size_t)1 << 40) / d) return COFHE_B200_EINVAL;
    cnt *= d;
  }
  if (bytes < 4 + 4 * (size_t)nd + 8 * cnt) return COFHE_B200_EINVAL;  // at least one table entry per element
  *ndim = nd;
  *n = cnt;
  return COFHE_B200_OK;
}

static int wire_slots_of(int kind) { return kind == 0 ? 6 : (kind == 1 ? 3 : 1); }

size_t cofhe_b200_wire_bound_kind(const cofhe_b200_ctx *ctx, int kind, uint32_t ndim, size_t n, uint32_t ml) {
  if (!ctx || kind < 0 || kind > 2) return 0;
  size_t per;
  if (kind == 2) per = 8 + (4 * (size_t)ml + 1);
  else per = (size_t)(kind == 0 ? 2 : 1) * (24 + 2 * (size_t)(4 * ctx->na + 1) + (size_t)(4 * ctx->nc + 1));
  return 4 + 4 * (size_t)ndim + n * per;
}
size_t cofhe_b200_wire_bound(const cofhe_b200_ctx *ctx, uint32_t ndim, size_t n) { return cofhe_b200_wire_bound_kind(ctx, 0, ndim, n, 0); }

int cofhe_b200_wire_to_batch(cofhe_b200_ctx *ctx, int kind, const uint8_t *wire, size_t bytes, uint32_t ndim, size_t n, uint32_t *out,
                             uint32_t ml, uint8_t *sign_out, void *stream) {
  if (!ctx || kind < 0 || kind > 2 || (n && (!wire || !out)) || (kind == 2 && (ml == 0 || ml > 4096))) return COFHE_B200_EINVAL;
  const size_t S = wire_slots_of(kind);
  size_t head = 4 + 4 * (size_t)ndim;
  if (bytes < head + 8 * S * n) return COFHE_B200_EINVAL;
  if (n == 0) return COFHE_B200_OK;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  WireShape s{(int)ctx->na, (int)ctx->nc, n, kind, (int)ml};
  unsigned blocks = (unsigned)((S * n + kWireBlock - 1) / kWireBlock);
  k_wire_read<<<blocks, kWireBlock, 0, (cudaStream_t)stream>>>(s, wire + head, wire + head + 8 * S * n,
                                                             (unsigned long long)(bytes - head - 8 * S * n), out, sign_out, ctx->d_err);
  return check_launch(ctx);
}
int cofhe_b200_wire_to_cts(cofhe_b200_ctx *ctx, const uint8_t *wire, size_t bytes, uint32_t ndim, size_t n, uint32_t *out,
                           void *stream) {
  return cofhe_b200_wire_to_batch(ctx, 0, wire, bytes, ndim, n, out, 0, nullptr, stream);
}

int cofhe_b200_batch_to_wire(cofhe_b200_ctx *ctx, int kind, uint32_t ndim, const uint32_t *dims, size_t n, const uint32_t *in, uint32_t ml,
                             const uint8_t *sign_in, uint8_t *wire, size_t cap, size_t *written, void *stream) {
  if (!ctx || kind < 0 || kind > 2 || !written || (ndim && !dims) || ndim > 8 || !wire || (n && !in) || (kind == 2 && (ml == 0 || ml > 4096)))
    return COFHE_B200_EINVAL;
  size_t cnt = 1;
  for (uint32_t i = 0; i < ndim; i++) cnt *= dims[i];
  if (cnt != n) return COFHE_B200_EINVAL;
  Enter g(ctx, (cudaStream_t)stream);
  if (!g.ok) return COFHE_B200_ECUDA;
  cudaStream_t st = (cudaStream_t)stream;
  const size_t S = wire_slots_of(kind);
  size_t head = 4 + 4 * (size_t)ndim;
  *written = head + 8 * S * n;
  if (cap < *written) return COFHE_B200_EINVAL;
  uint32_t hb[9];
  hb[0] = ndim;
  for (uint32_t i = 0; i < ndim; i++) hb[1 + i] = dims[i];
  CUDA_TRY(cudaMemcpyAsync(wire, hb, head, cudaMemcpyHostToDevice, st));
  CUDA_TRY(cudaStreamSynchronize(st));  // hb is a stack buffer
  if (n == 0) return COFHE_B200_OK;
  // scratch: offsets u64[m] | block totals u64[nb] | sum u64 | sizes u32[m] | flags u8[m]
  size_t m = S * n, nb = (m + kScanBlock - 1) / kScanBlock;
  int rc = ensure_work(ctx, (m + nb + 1) * 8 + m * 4 + m);
  if (rc != COFHE_B200_OK) return rc;
  unsigned long long *offs = (unsigned long long *)ctx->d_work, *totals = offs + m, *sum = totals + nb;
  u32 *sizes = (u32 *)(sum + 1);
  unsigned char *flags = (unsigned char *)(sizes + m);
  WireShape s{(int)ctx->na, (int)ctx->nc, n, kind, (int)ml};
  unsigned blocks = (unsigned)((m + kWireBlock - 1) / kWireBlock);
  k_wire_sizes<<<blocks, kWireBlock, 0, st>>>(s, in, sign_in, sizes, flags);
  ctx->launches++;
  k_scan_blocks<<<(unsigned)nb, kScanBlock, 0, st>>>(m, sizes, offs, totals);
  ctx->launches++;
  k_scan_totals<<<1, kScanBlock, 0, st>>>(nb, totals, sum);
  rc = check_launch(ctx);
  if (rc != COFHE_B200_OK) return rc;
  unsigned long long body = 0;
  CUDA_TRY(cudaMemcpyAsync(&body, sum, 8, cudaMemcpyDeviceToHost, st));
  CUDA_TRY(cudaStreamSynchronize(st));
  *written = head + 8 * S * n + (size_t)body;
  if (cap < *written) return COFHE_B200_EINVAL;  // *written tells the caller what is needed
  k_wire_write<<<blocks, kWireBlock, 0, st>>>(s, in, sizes, flags, offs, totals, wire + head, wire + head + 8 * S * n);
  return check_launch(ctx);
}
int cofhe_b200_cts_to_wire(cofhe_b200_ctx *ctx, uint32_t ndim, const uint32_t *dims, size_t n, const uint32_t *ct, uint8_t *wire,
                           size_t cap, size_t *written, void *stream) {
  return cofhe_b200_batch_to_wire(ctx, 0, ndim, dims, n, ct, 0, nullptr, wire, cap, written, stream);
}

}  // extern "C"

static double now_ms() {
  timespec ts;
  clock_gettime(CLOCK_MONOTONIC, &ts);
  return ts.tv_sec * 1e3 + ts.tv_nsec * 1e-6;
}

// Lazily created pipeline lane (stream + bail counters).
static int lane_ready(cofhe_b200_ctx *ctx, int l) {
  cofhe_b200_ctx::Lane &L = ctx->lane[l];
  if (L.st) return COFHE_B200_OK;
  if (cudaStreamCreateWithFlags(&L.st, cudaStreamNonBlocking) != cudaSuccess) { L.st = nullptr; return COFHE_B200_ECUDA; }
  if (cudaMalloc(&L.d_nbail, 2 * sizeof(int)) != cudaSuccess || cudaMemset(L.d_nbail, 0, 2 * sizeof(int)) != cudaSuccess)
    return COFHE_B200_ENOMEM;
  return COFHE_B200_OK;
}
// run `fn` with the lane's bail state installed in the context (launch_op reads it from there)
template <class F>
static int with_lane(cofhe_b200_ctx *ctx, int l, F fn) {
  cofhe_b200_ctx::Lane &L = ctx->lane[l];
  std::swap(ctx->d_bail, L.d_bail); std::swap(ctx->d_nbail, L.d_nbail);
  std::swap(ctx->bail_cap, L.bail_cap); std::swap(ctx->flip, L.flip);
  int rc = fn(L.st);
  std::swap(ctx->d_bail, L.d_bail); std::swap(ctx->d_nbail, L.d_nbail);
  std::swap(ctx->bail_cap, L.bail_cap); std::swap(ctx->flip, L.flip);
  return rc;
}

extern "C" {

/* Host-pointer ciphertext add: the batch is cut into element ranges that travel through up to 4 lanes, so the upload
 * of one range overlaps the kernels of the previous one and the download of the one before.  Host buffers are the
 * same SoA rows as everywhere ([rows][n]); a range is a strided 2-D copy into a dense [rows][cn] staging batch. */
int cofhe_b200_add_ciphertexts_host(cofhe_b200_ctx *ctx, size_t n, const uint32_t *ct1, const uint32_t *ct2,
                                    uint32_t *out) {
  if (!ctx || (n && (!ct1 || !ct2 || !out))) return COFHE_B200_EINVAL;
  if (ctx->compact) return COFHE_B200_ESTATE;  // compact variant: use upload + cofhe_b200_add_ciphertexts + download
  if (n == 0) return COFHE_B200_OK;
  EnterHost g(ctx);
  if (!g.ok) return COFHE_B200_ECUDA;
  const size_t rows = (size_t)2 * (2 * ctx->na + ctx->nc);
  size_t bytes = rows * n * sizeof(u32);
  int rc = COFHE_B200_OK;
  // Page-locked (cudaHostAlloc / cudaHostRegister) buffers: no staging at all.  The composition kernel reads the a and
  // b rows straight from host memory over PCIe (coalesced 128-byte requests; the c rows, 48 % of the bytes, are
  // redundant and recomputed on the device) and writes the result rows straight back, so the transfer of one warp
  // overlaps the arithmetic of the others inside a single launch.  Measured on B200 against copy-engine staging,
  // which pays a slow start after every idle gap: 0.29 ms vs 0.37-0.58 ms per 128x128 add.
  if (ctx->fastn && !ctx->opt.host_staged) {
    cudaPointerAttributes a1, a2, a3;
    bool pinned = cudaPointerGetAttributes(&a1, ct1) == cudaSuccess && cudaPointerGetAttributes(&a2, ct2) == cudaSuccess &&
                  cudaPointerGetAttributes(&a3, out) == cudaSuccess && a1.type == cudaMemoryTypeHost &&
                  a2.type == cudaMemoryTypeHost && a3.type == cudaMemoryTypeHost && a1.devicePointer && a2.devicePointer &&
                  a3.devicePointer;
    cudaGetLastError();
    if (pinned) {
      rc = lane_ready(ctx, 0);
      if (rc != COFHE_B200_OK) return rc;
      int flags = ctx->opt.host_loadc ? 0 : 2;  // 2: operands without c rows
      rc = with_lane(ctx, 0, [&](cudaStream_t s2) {
        return launch_op(ctx, 0, 2, n, (const u32 *)a1.devicePointer, (const u32 *)a2.devicePointer, (u32 *)a3.devicePointer, flags,
                         nullptr, 0, 0, nullptr, 0, s2);
      });
      if (rc == COFHE_B200_OK) rc = cofhe_b200_status(ctx, ctx->lane[0].st);
      else cudaStreamSynchronize(ctx->lane[0].st);
      return rc;
    }
  }
  if (ctx->stage_bytes < 3 * bytes) {
    cudaDeviceSynchronize();
    if (ctx->d_stage) cudaFree(ctx->d_stage);
    ctx->d_stage = nullptr; ctx->stage_bytes = 0;
    if (cudaMalloc(&ctx->d_stage, 3 * bytes) != cudaSuccess) { cudaGetLastError(); return COFHE_B200_ENOMEM; }
    ctx->stage_bytes = 3 * bytes;
  }
  // Default: two lanes, one per form block (all c1 forms, then all c2 forms: each block is contiguous in the SoA batch,
  // so its transfers are plain 1-D copies at full PCIe rate, and the c2 upload overlaps the c1 kernels).  Element
  // ranges (COFHE_B200_HOST_CHUNKS > 0) are strided 2-D copies, which the copy engine runs row by row (~1 us per
  // row): measured slower on B200 below ~10^5 elements.
  size_t chunks = ctx->opt.host_chunks > 0 ? (size_t)ctx->opt.host_chunks : 0;
  if (chunks == 0) {
    const size_t fw = (rows / 2) * n;  // words of one form block
    for (int y = 0; y < 2 && rc == COFHE_B200_OK; y++) {
      rc = lane_ready(ctx, y);
      if (rc != COFHE_B200_OK) break;
      cudaStream_t st = ctx->lane[y].st;
      u32 *d1 = ctx->d_stage + y * fw, *d2 = d1 + 2 * fw, *d3 = d2 + 2 * fw;
      if (cudaMemcpyAsync(d1, ct1 + y * fw, fw * 4, cudaMemcpyHostToDevice, st) != cudaSuccess ||
          cudaMemcpyAsync(d2, ct2 + y * fw, fw * 4, cudaMemcpyHostToDevice, st) != cudaSuccess) { rc = COFHE_B200_ECUDA; break; }
      rc = with_lane(ctx, y, [&](cudaStream_t s2) { return launch_op(ctx, 0, 1, n, d1, d2, d3, 0, nullptr, 0, 0, nullptr, 0, s2); });
      if (rc == COFHE_B200_OK && cudaMemcpyAsync(out + y * fw, d3, fw * 4, cudaMemcpyDeviceToHost, st) != cudaSuccess)
        rc = COFHE_B200_ECUDA;
    }
    if (cudaStreamSynchronize(ctx->lane[0].st) != cudaSuccess && rc == COFHE_B200_OK) rc = COFHE_B200_ECUDA;
    if (rc == COFHE_B200_OK) rc = cofhe_b200_status(ctx, ctx->lane[1].st);
    else if (ctx->lane[1].st) cudaStreamSynchronize(ctx->lane[1].st);
    return rc;
  }
  if (chunks > (n + 1023) / 1024) chunks = (n + 1023) / 1024;  // ranges below ~1k elements only add launch overhead
  size_t per = (((n + chunks - 1) / chunks) + 31) & ~(size_t)31;
  int used = 0;
  size_t c = 0;
  const bool trace = ctx->opt.host_trace != 0;  // debugging aid: device times of the first range
  cudaEvent_t ev[4];
  double t_start = 0;
  if (trace) { for (auto &e : ev) cudaEventCreate(&e); t_start = now_ms(); }
  for (size_t e0 = 0; e0 < n && rc == COFHE_B200_OK; e0 += per, c++) {
    size_t cn = n - e0 < per ? n - e0 : per;
    int l = (int)(c % 4);
    rc = lane_ready(ctx, l);
    if (rc != COFHE_B200_OK) break;
    if (l + 1 > used) used = l + 1;
    u32 *d1 = ctx->d_stage + 3 * rows * e0, *d2 = d1 + rows * cn, *d3 = d2 + rows * cn;
    cudaStream_t st = ctx->lane[l].st;
    if (trace && c == 0) cudaEventRecord(ev[0], st);
    if (cudaMemcpy2DAsync(d1, cn * 4, ct1 + e0, n * 4, cn * 4, rows, cudaMemcpyHostToDevice, st) != cudaSuccess ||
        cudaMemcpy2DAsync(d2, cn * 4, ct2 + e0, n * 4, cn * 4, rows, cudaMemcpyHostToDevice, st) != cudaSuccess) {
      rc = COFHE_B200_ECUDA;
      break;
    }
    if (trace && c == 0) cudaEventRecord(ev[1], st);
    rc = with_lane(ctx, l, [&](cudaStream_t s2) { return launch_op(ctx, 0, 2, cn, d1, d2, d3, 0, nullptr, 0, 0, nullptr, 0, s2); });
    if (trace && c == 0) cudaEventRecord(ev[2], st);
    if (rc == COFHE_B200_OK &&
        cudaMemcpy2DAsync(out + e0, n * 4, d3, cn * 4, cn * 4, rows, cudaMemcpyDeviceToHost, st) != cudaSuccess)
      rc = COFHE_B200_ECUDA;
    if (trace && c == 0) cudaEventRecord(ev[3], st);
  }
  double t_issue = trace ? now_ms() : 0;
  for (int l = 0; l < used; l++) {
    if (l + 1 < used) { if (cudaStreamSynchronize(ctx->lane[l].st) != cudaSuccess && rc == COFHE_B200_OK) rc = COFHE_B200_ECUDA; }
    else if (rc == COFHE_B200_OK) rc = cofhe_b200_status(ctx, ctx->lane[l].st);  // last lane: sync + sticky element status
    else cudaStreamSynchronize(ctx->lane[l].st);
  }
  if (trace) {
    double t_end = now_ms();
    float a = 0, b = 0, d = 0;
    cudaEventElapsedTime(&a, ev[0], ev[1]); cudaEventElapsedTime(&b, ev[1], ev[2]); cudaEventElapsedTime(&d, ev[2], ev[3]);
    fprintf(stderr, "add_host n=%zu chunks=%zu: issue %.3f ms, total %.3f ms; first range: H2D %.3f kernels %.3f D2H %.3f ms\n", n, c,
            t_issue - t_start, t_end - t_start, a, b, d);
    for (auto &e : ev) cudaEventDestroy(e);
  }
  return rc;
}

}  // extern "C"

// ---------------------------------------------------------------------------------------------------
// Integer-multiply roofline probe: independent 32x32->64 multiply-accumulate chains (IMAD.WIDE.U32), all SMs.
// bench.py divides algorithmic MACs/s by this measured rate (SURVEY.md section 8d "Governing roofline").
// ---------------------------------------------------------------------------------------------------
namespace {
template <int WIDE>
__global__ void __launch_bounds__(256) k_imad_probe(u32 *out, u32 seed, int iters) {
  u32 x = seed + threadIdx.x, y = seed * 7u + blockIdx.x;
  u64 a0 = x, a1 = y, a2 = x ^ y, a3 = x + y, a4 = x * 3u, a5 = y * 5u, a6 = x - y, a7 = ~x;
  u32 b0 = x, b1 = y, b2 = x ^ y, b3 = x + y, b4 = x * 3u, b5 = y * 5u, b6 = x - y, b7 = ~x;
  for (int i = 0; i < iters; i++) {
    if (WIDE) {
#define STEP(a) asm volatile("mad.wide.u32 %0, %1, %2, %0;" : "+l"(a) : "r"(x), "r"(y));
      STEP(a0) STEP(a1) STEP(a2) STEP(a3) STEP(a4) STEP(a5) STEP(a6) STEP(a7)
#undef STEP
    } else {
#define STEP(b) asm volatile("mad.lo.u32 %0, %1, %2, %0;" : "+r"(b) : "r"(x), "r"(y));
      STEP(b0) STEP(b1) STEP(b2) STEP(b3) STEP(b4) STEP(b5) STEP(b6) STEP(b7)
#undef STEP
    }
  }
  u64 s = a0 ^ a1 ^ a2 ^ a3 ^ a4 ^ a5 ^ a6 ^ a7;
  u32 t = b0 ^ b1 ^ b2 ^ b3 ^ b4 ^ b5 ^ b6 ^ b7;
  if (((u32)s ^ (u32)(s >> 32) ^ t) == 0x12345u) out[0] = 1;  // keep the chains alive
}
}  // namespace

extern "C" int cofhe_b200_imad_probe(int device, int wide, double *macs_per_s) {
  if (!macs_per_s) return COFHE_B200_EINVAL;
  int count = 0;
  if (cudaGetDeviceCount(&count) != cudaSuccess || device < 0 || device >= count) return COFHE_B200_ECUDA;
  DevGuard g(device);
  if (!g.ok) return COFHE_B200_ECUDA;
  cudaDeviceProp prop;
  CUDA_TRY(cudaGetDeviceProperties(&prop, device));
  u32 *d = nullptr;
  CUDA_TRY(cudaMalloc(&d, 4));
  const int iters = 1 << 14, blocks = prop.multiProcessorCount * 8, threads = 256;
  cudaEvent_t e0, e1;
  CUDA_TRY(cudaEventCreate(&e0));
  CUDA_TRY(cudaEventCreate(&e1));
  double best = 0;
  for (int rep = 0; rep < 6; rep++) {
    cudaEventRecord(e0);
    if (wide) k_imad_probe<1><<<blocks, threads>>>(d, 12345u + rep, iters);
    else k_imad_probe<0><<<blocks, threads>>>(d, 12345u + rep, iters);
    cudaEventRecord(e1);
    if (cudaEventSynchronize(e1) != cudaSuccess) { cudaFree(d); return COFHE_B200_ECUDA; }
    float ms = 0;
    cudaEventElapsedTime(&ms, e0, e1);
    double rate = (double)blocks * threads * iters * 8.0 / (ms * 1e-3);
    if (rep > 0 && rate > best) best = rate;
  }
  cudaEventDestroy(e0);
  cudaEventDestroy(e1);
  cudaFree(d);
  *macs_per_s = best;
  return COFHE_B200_OK;
}
