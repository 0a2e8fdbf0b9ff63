// cofhe_b200 -- register-resident kernels for 4-limb coefficients (robust-layer capacity 32): see fast_launch.cuh.
#include "fast_launch_impl.cuh"

template struct cofhe::FastKernels<4, 32>;
