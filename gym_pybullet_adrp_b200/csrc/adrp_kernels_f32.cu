// fp32 product build of every kernel.
#include "adrp_kernels_impl.cuh"
template struct adrp::Launch<float>;
