// fp32 product build of the race step kernel for Physics.PYB (mask ADRP_PHY_GROUND = 16): BASELINE config C3.
#include "adrp_race_special.cuh"
template adrp::RaceStepFnF32 adrp::race_step_special<16>(int, bool);
