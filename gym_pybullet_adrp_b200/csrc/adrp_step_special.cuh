// adrp_step_special.cuh -- step_kernel instances of the fp32 product build for ONE physics mask (template argument), RPM-type
// actions, 1 / 2 / 4 / 8 lanes per env, with the env family compiled in (CtrlAviary-kind or RL aviary).  Included by
// adrp_step_f32_plain.cu / adrp_step_f32_forces.cu only.
#pragma once
#include "adrp_kernels.cuh"
#include "adrp_step.cuh"

namespace adrp {

template <int PHY, int ENVK, bool MULTI>
StepFnF32 step_special_mp(int mpad)
{
    switch (mpad) {
        case 1: return step_kernel<float, PHY, 0, 1, ENVK, MULTI, 0>;
        case 2: return step_kernel<float, PHY, 0, 2, ENVK, MULTI, 0>;
        case 4: return step_kernel<float, PHY, 0, 4, ENVK, MULTI, 0>;
        case 8: return step_kernel<float, PHY, 0, 8, ENVK, MULTI, 0>;
        default: return nullptr;
    }
}

// (instances for the X prop layout -- CF2X, RACE; a CF2P handle runs the generic kernel)
// ctrl: CtrlAviary-kind env; rl_too: this mask also has RL-aviary instances (the plain masks: learn.py / HoverAviary shapes)
template <int PHY, bool MULTI>
StepFnF32 step_special_impl(int mpad, bool ctrl)
{
    constexpr bool rl_too = PHY == ADRP_PHY_DYN || PHY == ADRP_PHY_GROUND;
    if (ctrl) return step_special_mp<PHY, 0, MULTI>(mpad);
    if constexpr (rl_too) return step_special_mp<PHY, 1, MULTI>(mpad);
    return nullptr;
}
#ifdef ADRP_MULTI_TU
// envk: 0 = CtrlAviary-kind, 1 = RL aviary, 2 = CtrlAviary-kind with the 12-column observation row (adrp_config.obs_columns)
// compiled in -- its own instances, because the run-time choice of the row moves the allocation of the 20-column ones (step_kernel)
template <int PHY> StepFnF32 step_special_multi(int mpad, int envk)
{
    if (envk == 2) return step_special_mp<PHY, 2, true>(mpad);
    return step_special_impl<PHY, true>(mpad, envk == 0);
}
#else
template <int PHY> StepFnF32 step_special(int mpad, bool ctrl) { return step_special_impl<PHY, false>(mpad, ctrl); }
#endif

}  // namespace adrp
