// cooperative kernels: 2 lanes x 5 limbs per form, 7-limb coefficients
#include "coop_launch_impl.cuh"
namespace cofhe { template struct CoopKernels<2, 5, 7, 32>; }
