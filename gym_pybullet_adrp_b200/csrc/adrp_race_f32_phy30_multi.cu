// fp32 product build of the race step kernel for Physics.PYB_GND_DRAG_DW, adrp_step_n instances (K control steps per launch).
#define ADRP_MULTI_TU
#include "adrp_race_special.cuh"
template adrp::RaceStepFnF32 adrp::race_step_special_multi<30>(int, bool);
