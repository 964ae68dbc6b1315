// adrp_step_special.cuh -- step_kernel instances of the fp32 product build for ONE physics mask (template argument), RPM-type
// actions, 1 / 2 / 4 / 8 lanes per env.  Included by adrp_step_f32_plain.cu / adrp_step_f32_forces.cu only.
#pragma once
#include "adrp_kernels.cuh"
#include "adrp_step.cuh"

namespace adrp {

template <int PHY>
StepFnF32 step_special(int mpad)
{
    switch (mpad) {
        case 1: return step_kernel<float, PHY, 0, 1>;
        case 2: return step_kernel<float, PHY, 0, 2>;
        case 4: return step_kernel<float, PHY, 0, 4>;
        case 8: return step_kernel<float, PHY, 0, 8>;
        default: return nullptr;
    }
}

}  // namespace adrp
