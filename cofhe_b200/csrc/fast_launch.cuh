// cofhe_b200 -- launch interface of the register-resident kernels, one explicit instantiation per limb count.
//
// The fixed-width kernels are large (a composition is ~25 k SASS instructions per instantiation), so each limb count
// N lives in its own translation unit (fast_n<N>.cu: `template struct FastKernels<N, CAP>`), compiled in parallel;
// kernels.cu sees only these declarations.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

#include "matmul_shape.cuh"

namespace cofhe {

constexpr int kBlock = 128;
// Exponentiation and matmul kernels are compiled for up to 512 threads per block (128 registers: 16 warps per SM, all
// of them aligned by the block barriers) and launched with 128, 256 or 512: a batch that cannot fill the machine is
// bound by the latency of one warp, which is shortest with few warps per SM on as many SMs as possible.
constexpr int kPowBlock = 512;
constexpr int kMMBlock = 512;
// threads per block for `threads` independent threads on `sms` multiprocessors: the largest block that still gives
// every SM one
inline int pick_block(size_t threads, int sms) {
  if (sms <= 0) sms = 148;
  if (threads >= (size_t)sms * 512) return 512;
  if (threads >= (size_t)sms * 256) return 256;
  return 128;
}

template <int FN, int FCAP>
struct FastKernels {
  // G: const Group<FCAP>* in device memory.  flags of compose: 1 = second operand inverted, 2 = operands without c rows
  static void compose(bool dupl, dim3 grid, cudaStream_t st, const void *G, int dbits, int nc, size_t n, const u32 *in1,
                      const u32 *in2, u32 *out, int flags, u32 *bail, int *nbail);
  static void pow(dim3 grid, int block, cudaStream_t st, const void *G, int dbits, int nc, size_t n, const u32 *in, u32 *out, const u32 *exp,
                  int el, int per_elem, const unsigned char *expneg, int shared_neg, u32 *bail, int *nbail);
  static void mm_chain(dim3 grid, int block, cudaStream_t st, const void *G, int dbits, int nc, MMShape s, const u32 *ct, u32 *tab);
  static void mm_fill(dim3 grid, int block, cudaStream_t st, const void *G, int dbits, MMShape s, u32 *tab);
  static void mm_accum(dim3 grid, int block, cudaStream_t st, const void *G, int dbits, MMShape s, const u32 *scal, const u32 *tab, u32 *part);
  static void mm_final(dim3 grid, int block, cudaStream_t st, const void *G, int dbits, int nc, MMShape s, const u32 *init, const u32 *part,
                       u32 *out, u32 *bail, int *nbail);
};

}  // namespace cofhe
