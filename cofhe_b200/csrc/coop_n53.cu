// cooperative kernels: 8 lanes x 7 limbs per form, 53-limb coefficients (3333-bit discriminants: N of 3072 bits, k = 128)
#include "coop_launch_impl.cuh"
namespace cofhe { template struct CoopKernels<8, 7, 53, 216>; }
