// fp64 parity build of every kernel (compiled with -fmad=false, see build.py).
#include "adrp_kernels_impl.cuh"
template struct adrp::Launch<double>;
