// adrp_kernels.cuh -- launch interface between the host side of the C ABI (adrp_api.cu) and the two
// kernel translation units: adrp_kernels_f32.cu (fp32 product build) and adrp_kernels_f64.cu (fp64
// parity build, compiled with -fmad=false so that its arithmetic is statement-for-statement IEEE).
#pragma once
#include <cuda_runtime.h>

#include "adrp_params.cuh"

namespace adrp {

template <typename T>
struct Launch {
    static cudaError_t reset(const KParams<T>& P, const KBuffers<T>& b, const uint8_t* mask, T* obs, cudaStream_t s);
    static cudaError_t step(const KParams<T>& P, const KBuffers<T>& b, const T* action, T* obs, T* reward,
                            uint8_t* terminated, uint8_t* truncated, uint8_t* flags, cudaStream_t s);
    static cudaError_t set_state(const KParams<T>& P, const KBuffers<T>& b, const T* rows, cudaStream_t s);
    static cudaError_t get_state(const KParams<T>& P, const KBuffers<T>& b, T* rows, cudaStream_t s);
    static cudaError_t race_prepare(const RaceParams<T>& RP, const KParams<T>& P);     // shared-memory opt-in
    static cudaError_t race_reset(const KParams<T>& P, const RaceParams<T>& RP, const KBuffers<T>& b, const RaceBuffers<T>& rb,
                                  const uint8_t* mask, const T* draws, T* obs, cudaStream_t s);
    static cudaError_t race_step(const KParams<T>& P, const RaceParams<T>& RP, const KBuffers<T>& b, const RaceBuffers<T>& rb,
                                 const T* action, const T* noise, T* obs, T* reward, uint8_t* terminated,
                                 uint8_t* truncated, uint8_t* flags, cudaStream_t s);
    static cudaError_t dslpid_rows(T pid_gravity, T pid_kf, int model, int n, T dt, const T* pos, const T* quat, const T* vel,
                                   const T* tpos, const T* trpy, const T* tvel, const T* trates, T* ctrl, T* rpm, T* pos_e,
                                   T* yaw_e, cudaStream_t s);
};

// Race step kernels of the product build with the physics mask as a template argument, one translation unit per mask
// (adrp_race_f32_phy16.cu: Physics.PYB, adrp_race_f32_phy30.cu: Physics.PYB_GND_DRAG_DW).  NULL: no such instance.
using RaceStepFnF32 = void (*)(KParams<float>, RaceParams<float>, KBuffers<float>, RaceBuffers<float>, const float*, const float*,
                               float*, float*, uint8_t*, uint8_t*, uint8_t*);
template <int PHY> RaceStepFnF32 race_step_special(int mpad, bool quiet);
template <int PHY> RaceStepFnF32 race_step_special_multi(int mpad, bool quiet);   // adrp_step_n: K control steps per launch
template <int PHY> RaceStepFnF32 race_step_special_dense(int mpad, bool quiet, bool multi);   // 128-register instances (multi-wave grids)

// Likewise for step_kernel (RPM-type actions): the plain masks (Physics.DYN = 1, Physics.PYB = 16; adrp_step_f32_plain.cu) and the
// all-forces masks (DYN + GND + DRAG + DW = 15, Physics.PYB_GND_DRAG_DW = 30; adrp_step_f32_forces.cu), 1 / 2 / 4 / 8 lanes per env.
using StepFnF32 = void (*)(KParams<float>, KBuffers<float>, const float*, float*, float*, uint8_t*, uint8_t*, uint8_t*);
template <int PHY> StepFnF32 step_special(int mpad, bool ctrl);      // NULL: no such instance
template <int PHY> StepFnF32 step_special_multi(int mpad, int envk);    // adrp_step_n: K control steps per launch (envk: adrp_step_special.cuh)

}  // namespace adrp
