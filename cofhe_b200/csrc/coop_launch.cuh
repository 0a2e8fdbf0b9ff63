// cofhe_b200 -- launch interface of the cooperative kernels (coop_launch_impl.cuh), one explicit instantiation per
// (lanes per form T, limbs per lane L, coefficient limbs N, capacity class of the robust layer) in its own translation
// unit coop_n<N>.cu; kernels.cu sees only these declarations.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

#include "fw.cuh"
#include "matmul_shape.cuh"

namespace cofhe {

constexpr int kCoopBlock = 128;

template <int T, int L, int N, int CAP>
struct CoopKernels {
  // flags of compose: 1 = second operand inverted
  static void compose(bool dupl, unsigned ny, cudaStream_t st, const void *G, int dbits, int nc, size_t n, const u32 *in1, const u32 *in2,
                      u32 *out, int flags, u32 *bail, int *nbail, int mb);
  // out[e] = in[e]^x (x shared: exp[0..el), or per element: exp[l*n + e]; sign applied at the end); tab: scratch of
  // pow_tab_words(ny, nc, n) words for the window tables
  static void pow(unsigned ny, cudaStream_t st, const void *G, int dbits, int nc, size_t n, const u32 *in, u32 *out, const u32 *exp, int el, int per_elem,
                  const unsigned char *expneg, int shared_neg, u32 *tab, u32 *bail, int *nbail, int block);
  static size_t pow_tab_words(unsigned ny, int nc, size_t n);
  // multi-exponentiation from SoA window tables (coop_ops.cuh cq_tabprod): part[(chunk*ny + y)][fl][n*p]; outputs to redo
  // are appended to bail as y * n*p + index
  static void tabprod(unsigned ny, cudaStream_t st, const void *G, int dbits, int nc, MMShape s, const u32 *scal, const u32 *tab, size_t ne,
                      const u32 *init, u32 *part, u32 *bail, int *nbail, int block);
};

}  // namespace cofhe
