// fp32 product build of the race step kernel for Physics.PYB_GND_DRAG_DW (mask GND | DRAG | DW | GROUND = 30): the
// north_star's target configuration (C5M).
#include "adrp_race_special.cuh"
template adrp::RaceStepFnF32 adrp::race_step_special<30>(int, bool);
