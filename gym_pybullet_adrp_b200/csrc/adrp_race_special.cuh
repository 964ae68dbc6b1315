// adrp_race_special.cuh -- race step kernels of the fp32 product build for ONE physics mask (template argument), lanes
// per env 1 / 2 / 4 / 8, with and without disturbance draws.  Included by adrp_race_f32_phy<mask>.cu only.
#pragma once
#include "adrp_kernels.cuh"
#include "adrp_race.cuh"

namespace adrp {

template <int PHY>
RaceStepFnF32 race_step_special(int mpad, bool quiet)
{
    switch (mpad) {
        case 1: return quiet ? race_step_kernel<float, 1, false, true, PHY> : race_step_kernel<float, 1, false, false, PHY>;
        case 2: return quiet ? race_step_kernel<float, 2, false, true, PHY> : race_step_kernel<float, 2, false, false, PHY>;
        case 4: return quiet ? race_step_kernel<float, 4, false, true, PHY> : race_step_kernel<float, 4, false, false, PHY>;
        case 8: return quiet ? race_step_kernel<float, 8, false, true, PHY> : race_step_kernel<float, 8, false, false, PHY>;
        default: return nullptr;
    }
}

}  // namespace adrp
