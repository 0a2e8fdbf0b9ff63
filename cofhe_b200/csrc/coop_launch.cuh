// cofhe_b200 -- launch interface of the cooperative kernels (coop_launch_impl.cuh), one explicit instantiation per
// (lanes per form T, limbs per lane L, coefficient limbs N, capacity class of the robust layer) in its own translation
// unit coop_n<N>.cu; kernels.cu sees only these declarations.
#pragma once
#include <cuda_runtime.h>
#include <stddef.h>

#include "fw.cuh"

namespace cofhe {

constexpr int kCoopBlock = 128;

template <int T, int L, int N, int CAP>
struct CoopKernels {
  // flags of compose: 1 = second operand inverted
  static void compose(bool dupl, unsigned ny, cudaStream_t st, const void *G, int dbits, int nc, size_t n, const u32 *in1, const u32 *in2,
                      u32 *out, int flags, u32 *bail, int *nbail, int mb);
};

}  // namespace cofhe
