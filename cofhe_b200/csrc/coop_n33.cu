// cooperative kernels: 4 lanes x 9 limbs per form, 33-limb coefficients
#include "coop_launch_impl.cuh"
namespace cofhe { template struct CoopKernels<4, 9, 33, 136>; }
