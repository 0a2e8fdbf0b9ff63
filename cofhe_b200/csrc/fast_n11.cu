// cofhe_b200 -- register-resident kernels for 11-limb coefficients (robust-layer capacity 48): see fast_launch.cuh.
#include "fast_launch_impl.cuh"

template struct cofhe::FastKernels<11, 48>;
