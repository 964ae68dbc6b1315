// fp32 product build of step_kernel, adrp_step_n instances (K control steps per launch) of the all-forces masks.
#define ADRP_MULTI_TU
#include "adrp_step_special.cuh"
template adrp::StepFnF32 adrp::step_special_multi<15>(int, int);
template adrp::StepFnF32 adrp::step_special_multi<30>(int, int);
