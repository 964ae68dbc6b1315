// adrp_race_special.cuh -- race step kernels of the fp32 product build for ONE physics mask (template argument), lanes
// per env 1 / 2 / 4 / 8, with and without disturbance draws.  Included by adrp_race_f32_phy<mask>.cu only.
#pragma once
#include "adrp_kernels.cuh"
#include "adrp_race.cuh"

namespace adrp {

template <int PHY, bool MULTI, int MINB>
RaceStepFnF32 race_step_special_impl(int mpad, bool quiet)
{
    switch (mpad) {
        case 1: return quiet ? race_step_kernel<float, 1, 0, true, PHY, MULTI, MINB> : race_step_kernel<float, 1, 0, false, PHY, MULTI, MINB>;
        case 2: return quiet ? race_step_kernel<float, 2, 0, true, PHY, MULTI, MINB> : race_step_kernel<float, 2, 0, false, PHY, MULTI, MINB>;
        case 4: return quiet ? race_step_kernel<float, 4, 0, true, PHY, MULTI, MINB> : race_step_kernel<float, 4, 0, false, PHY, MULTI, MINB>;
        case 8: return quiet ? race_step_kernel<float, 8, 0, true, PHY, MULTI, MINB> : race_step_kernel<float, 8, 0, false, PHY, MULTI, MINB>;
        default: return nullptr;
    }
}
#if defined(ADRP_DENSE_TU)
template <int PHY> RaceStepFnF32 race_step_special_dense(int mpad, bool quiet, bool multi)
{
    return multi ? race_step_special_impl<PHY, true, kRaceMinBlocksDense>(mpad, quiet) : race_step_special_impl<PHY, false, kRaceMinBlocksDense>(mpad, quiet);
}
#elif defined(ADRP_MULTI_TU)
template <int PHY> RaceStepFnF32 race_step_special_multi(int mpad, bool quiet) { return race_step_special_impl<PHY, true, kRaceMinBlocks>(mpad, quiet); }
#else
template <int PHY> RaceStepFnF32 race_step_special(int mpad, bool quiet) { return race_step_special_impl<PHY, false, kRaceMinBlocks>(mpad, quiet); }
#endif

}  // namespace adrp
