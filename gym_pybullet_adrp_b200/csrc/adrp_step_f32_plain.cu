// fp32 product build of step_kernel for the masks without optional force models: Physics.DYN (1) and Physics.PYB (16, BASELINE
// config C2 and the learn.py shape).
#include "adrp_step_special.cuh"
template adrp::StepFnF32 adrp::step_special<1>(int, bool);
template adrp::StepFnF32 adrp::step_special<16>(int, bool);
