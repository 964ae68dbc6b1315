// fp32 product build of step_kernel for the masks with every force model on: DYN + GND + DRAG + DW (15, BASELINE config C5) and
// Physics.PYB_GND_DRAG_DW (30, config C4).
#include "adrp_step_special.cuh"
template adrp::StepFnF32 adrp::step_special<15>(int, bool);
template adrp::StepFnF32 adrp::step_special<30>(int, bool);
