// fp32 product build of the race step kernel for Physics.PYB, DENSE instances (128 registers: grids longer than one wave),
// single-step and adrp_step_n.
#define ADRP_DENSE_TU
#include "adrp_race_special.cuh"
template adrp::RaceStepFnF32 adrp::race_step_special_dense<16>(int, bool, bool);
