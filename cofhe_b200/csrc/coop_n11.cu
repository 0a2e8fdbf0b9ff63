// cooperative kernels: 2 lanes x 7 limbs per form, 11-limb coefficients
#include "coop_launch_impl.cuh"
namespace cofhe { template struct CoopKernels<2, 7, 11, 48>; }
