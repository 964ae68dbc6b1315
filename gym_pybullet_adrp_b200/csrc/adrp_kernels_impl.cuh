// adrp_kernels_impl.cuh -- Launch<T> bodies; included by exactly one .cu per precision.
#pragma once
#include <cstdlib>
#include <map>
#include <mutex>
#include <utility>

#include "adrp_kernels.cuh"
#include "adrp_race.cuh"

namespace adrp {

static_assert(kBlock == 128, "adrp_api.cu cuts host-path chunks in units of 128-thread blocks");

// Launch `threads` threads of `kernel`.
// Block size: grids longer than one wave of resident blocks use block_max.  A grid that fits one wave is cut into the
// block size (block_max, block_max / 2, ... >= 32) that loads the busiest SM least: BASELINE's C2 is 65 536 threads = 512
// blocks of 128 on 148 SMs -- 3.46 per SM, i.e. SMs with 4 blocks set the time (512 threads) -- but 1 024 blocks of 64 are
// 7 per SM at most (448 threads); C3's 32 768 threads are 4 x 64 vs 7 x 32.  The kernels index with blockDim.x.
// ADRP_BLOCK=<n> forces a size (diagnostic).
// Programmatic stream serialisation (see pdl_wait in adrp_step.cuh) when the grid is longer than one wave: the next
// step's blocks then fill the slots the last wave frees (C4 +13 %, C5 +2 %).  A grid that fits one wave, launched that way as
// it is, has its dependents placed into whatever slots the running grid leaves, unevenly over the SMs (measured: C2 -17 %,
// C3 -13 %).  Single-step launches of such a grid are therefore padded with dynamic shared memory until an SM holds exactly
// ceil(blocks / SMs) blocks: the dependents can only take the slots of blocks that exit, spread like them, and wait at
// griddepcontrol.wait already resident when the grid before them drains -- the launch gap is gone (C2 4.56e9 -> 5.65e9, C3
// +3 %, C1 +2 %).  The padding takes L1 capacity, which the long n-step launches (no gap to hide) would only pay for
// (C2 adrp_step_n -1.7 %): they launch plainly.  ADRP_PDL_ONEWAVE=0 switches this off, ADRP_PDL=0 every programmatic launch.
constexpr size_t kPadBudget = 200 * 1024;   // dynamic shared memory per SM handed out as padding (of 227 KB usable)
template <typename... KArgs, typename... Args>
inline void launch_pdl(void (*kernel)(KArgs...), long long threads, unsigned block_max, size_t smem_at_max, cudaStream_t s, bool single_step,
                       Args... args)
{
    static const bool enabled = [] { const char* e = getenv("ADRP_PDL"); return !e || e[0] != '0'; }();
    static const int forced = [] { const char* e = getenv("ADRP_BLOCK"); return e ? atoi(e) : 0; }();
    static const bool onewave_pdl = [] { const char* e = getenv("ADRP_PDL_ONEWAVE"); return !e || e[0] != '0'; }();
    static std::mutex mu;
    struct Info { long long wave; int sms; bool pad_ok; };
    static std::map<std::pair<const void*, int>, Info> cache;     // (kernel, device) -> resident blocks at block_max, SMs
    Info info;
    {
        int dev = 0;
        cudaGetDevice(&dev);
        std::lock_guard<std::mutex> lock(mu);
        auto key = std::make_pair((const void*)kernel, dev);
        auto it = cache.find(key);
        if (it == cache.end()) {
            int per_sm = 0, sms = 0;
            cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, kernel, (int)block_max, smem_at_max);
            // (the padding below; a device that cannot give a block that much shared memory launches one-wave grids plainly)
            const bool pad_ok = cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)kPadBudget) == cudaSuccess;
            if (!pad_ok) cudaGetLastError();
            cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
            it = cache.emplace(key, Info{(long long)per_sm * sms, sms, pad_ok}).first;
        }
        info = it->second;
    }
    unsigned block = block_max;
    const long long grid_max = (threads + block_max - 1) / block_max;
    const bool one_wave = info.wave > 0 && grid_max <= info.wave;
    if (forced >= 32 && forced <= (int)block_max && block_max % forced == 0) block = (unsigned)forced;
    else if (one_wave && info.sms > 0) {
        long long best = -1;
        for (unsigned b = block_max; b >= 32; b >>= 1) {
            const long long blocks = (threads + b - 1) / b;
            const long long busiest = (blocks + info.sms - 1) / info.sms * b;      // threads on the fullest SM
            if (best < 0 || busiest < best) { best = busiest; block = b; }
        }
    }
    cudaLaunchConfig_t cfg = {};
    cfg.gridDim = dim3((unsigned)((threads + block - 1) / block), 1, 1);
    cfg.blockDim = dim3(block, 1, 1);
    cfg.dynamicSmemBytes = smem_at_max * block / block_max;
    cfg.stream = s;
    bool pdl = enabled && !one_wave;
    if (enabled && one_wave && onewave_pdl && single_step && info.pad_ok && info.sms > 0) {
        const long long per_sm = (cfg.gridDim.x + info.sms - 1) / info.sms;
        const size_t pad = (kPadBudget / (size_t)per_sm) & ~(size_t)1023;
        if (pad > cfg.dynamicSmemBytes) { cfg.dynamicSmemBytes = pad; pdl = true; }
    }
    cudaLaunchAttribute attr[1];
    attr[0].id = cudaLaunchAttributeProgrammaticStreamSerialization;
    attr[0].val.programmaticStreamSerializationAllowed = 1;
    cfg.attrs = attr;
    cfg.numAttrs = pdl ? 1 : 0;
    cudaLaunchKernelEx(&cfg, kernel, KArgs(args)...);
}

inline unsigned grid_for(long long D) { return (unsigned)((D + kBlock - 1) / kBlock); }

template <typename T>
inline size_t race_smem_bytes(const KParams<T>& P, const RaceParams<T>& RP, int block = kBlock) { return sizeof(T) * (size_t)RP.TF * (block / P.Mpad); }

// DSLPIDControl.computeControl for n independent rows (control/DSLPIDControl.py:82-145).
template <typename T>
__global__ void dslpid_rows_kernel(T pid_gravity, T pid_kf, int model, int n, T dt, const T* pos, const T* quat,
                                   const T* vel, const T* tpos, const T* trpy, const T* tvel, const T* trates,
                                   T* ctrl, T* rpm, T* pos_e, T* yaw_e)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    PidState<T> c;
    T* cs = ctrl + 9 * r;
    c.ipos = {cs[0], cs[1], cs[2]}; c.irpy = {cs[3], cs[4], cs[5]}; c.last_rpy = {cs[6], cs[7], cs[8]};
    const V3<T> zero = {T(0), T(0), T(0)};
    auto v3 = [&](const T* p) -> V3<T> { return p ? V3<T>{p[3 * r], p[3 * r + 1], p[3 * r + 2]} : zero; };
    const Q4<T> q = {quat[4 * r], quat[4 * r + 1], quat[4 * r + 2], quat[4 * r + 3]};
    T out[4], ye;
    V3<T> pe;
    dslpid_control(pid_gravity, pid_kf, model, dt, v3(pos), q, v3(vel), v3(tpos), v3(trpy), v3(tvel), v3(trates), c, out, pe, ye);
    for (int k = 0; k < 4; ++k) rpm[4 * r + k] = out[k];
    if (pos_e) { pos_e[3 * r] = pe.x; pos_e[3 * r + 1] = pe.y; pos_e[3 * r + 2] = pe.z; }
    if (yaw_e) yaw_e[r] = ye;
    cs[0] = c.ipos.x; cs[1] = c.ipos.y; cs[2] = c.ipos.z; cs[3] = c.irpy.x; cs[4] = c.irpy.y; cs[5] = c.irpy.z;
    cs[6] = c.last_rpy.x; cs[7] = c.last_rpy.y; cs[8] = c.last_rpy.z;
}

template <typename T>
cudaError_t Launch<T>::reset(const KParams<T>& P, const KBuffers<T>& b, const uint8_t* mask, T* obs, cudaStream_t s)
{
    reset_kernel<T><<<grid_for(P.Dpad), kBlock, 0, s>>>(P, b, mask, obs);
    return cudaGetLastError();
}

template <typename T>
cudaError_t Launch<T>::step(const KParams<T>& P, const KBuffers<T>& b, const T* action, T* obs, T* reward,
                            uint8_t* te, uint8_t* tr, uint8_t* fl, cudaStream_t s)
{
    const bool pid = P.action_type == ADRP_ACT_PID || P.action_type == ADRP_ACT_ONE_D_PID ||
                     P.action_type == ADRP_ACT_VEL || P.action_type == ADRP_ACT_PID_FULL;
    const bool multi = P.K > 1 || P.act_rng;                  // adrp_step_n: the instances with the control-step loop (and the action draw)
    const long long grid = (long long)(P.env_end - P.env_begin) * P.Mpad;      // threads of this launch
    // product build: the masks the configs use (Physics.DYN, Physics.PYB and the two all-forces masks) have instances with the
    // mask as a template argument (no per-substep mask tests, fewer registers: C5 +2 %, C2 +2 % at every batch size measured),
    // compiled in their own translation units (step_special)
    if constexpr (M<T>::fast) {
        static const bool generic = [] { const char* e = getenv("ADRP_STEP_GENERIC"); return e && e[0] == '1'; }();   // diagnostic
        // (the instances have the X prop layout compiled in; the CtrlAviary-kind n-step ones lack the in-kernel action draw)
        if (!pid && !generic && P.drone_model != ADRP_CF2P && !(P.act_rng && P.env_kind == ADRP_ENV_CTRL)) {
            StepFnF32 f = nullptr;
            const bool ctrl = P.env_kind == ADRP_ENV_CTRL;
            const int envk = ctrl ? (P.O == 12 ? 2 : 0) : 1;
            constexpr int ALL = ADRP_PHY_GND | ADRP_PHY_DRAG | ADRP_PHY_DW;
            switch (P.physics) {
                case ADRP_PHY_DYN: f = multi ? step_special_multi<ADRP_PHY_DYN>(P.Mpad, envk) : step_special<ADRP_PHY_DYN>(P.Mpad, ctrl); break;
                case ADRP_PHY_GROUND: f = multi ? step_special_multi<ADRP_PHY_GROUND>(P.Mpad, envk) : step_special<ADRP_PHY_GROUND>(P.Mpad, ctrl); break;
                case ADRP_PHY_DYN | ALL: f = multi ? step_special_multi<ADRP_PHY_DYN | ALL>(P.Mpad, envk) : step_special<ADRP_PHY_DYN | ALL>(P.Mpad, ctrl); break;
                case ADRP_PHY_GROUND | ALL: f = multi ? step_special_multi<ADRP_PHY_GROUND | ALL>(P.Mpad, envk) : step_special<ADRP_PHY_GROUND | ALL>(P.Mpad, ctrl); break;
                default: break;
            }
            if (f) { launch_pdl(f, grid, kStepBlock, 0, s, !multi, P, b, action, obs, reward, te, tr, fl); return cudaGetLastError(); }
        }
    }
#define ADRP_LAUNCH_STEP(MP) launch_pdl(step_kernel<T, -1, 0, MP>, grid, kStepBlock, 0, s, true, P, b, action, obs, reward, te, tr, fl)
    if (multi) {                                              // generic instances with the control-step loop
        if (pid) launch_pdl(step_kernel<T, -1, 1, 0, -1, true>, grid, kStepBlock, 0, s, false, P, b, action, obs, reward, te, tr, fl);
        else launch_pdl(step_kernel<T, -1, 0, 0, -1, true>, grid, kStepBlock, 0, s, false, P, b, action, obs, reward, te, tr, fl);
    } else if (pid) launch_pdl(step_kernel<T, -1, 1, 0>, grid, kStepBlock, 0, s, true, P, b, action, obs, reward, te, tr, fl);
    else switch (P.Mpad) {
        case 1: ADRP_LAUNCH_STEP(1); break;
        case 2: ADRP_LAUNCH_STEP(2); break;
        case 4: ADRP_LAUNCH_STEP(4); break;
        case 8: ADRP_LAUNCH_STEP(8); break;
        case 16: ADRP_LAUNCH_STEP(16); break;
        default: ADRP_LAUNCH_STEP(32); break;
    }
#undef ADRP_LAUNCH_STEP
    return cudaGetLastError();
}

template <typename T>
cudaError_t Launch<T>::set_state(const KParams<T>& P, const KBuffers<T>& b, const T* rows, cudaStream_t s)
{
    set_state_kernel<T><<<grid_for(P.Dpad), kBlock, 0, s>>>(P, b, rows);
    return cudaGetLastError();
}

template <typename T>
cudaError_t Launch<T>::get_state(const KParams<T>& P, const KBuffers<T>& b, T* rows, cudaStream_t s)
{
    get_state_kernel<T><<<grid_for(P.Dpad), kBlock, 0, s>>>(P, b, rows);
    return cudaGetLastError();
}

// The race step kernel the launch uses: the product build unrolls the neighbour loops for 1 / 2 / 4 / 8 lanes per
// env; the parity build (and wider envs) read the lane count from the params.
template <typename T>
using RaceStepFn = void (*)(KParams<T>, RaceParams<T>, KBuffers<T>, RaceBuffers<T>, const T*, const T*, T*, T*, uint8_t*,
                            uint8_t*, uint8_t*);
// quiet: no disturbance draws in this launch (no injected noise, RaceParams::disturbances off) -- the product build
// then runs the instance without the wind / motor-noise terms; phy: the physics mask, for which the product build has
// instances in their own translation units (race_step_special).
template <typename T>
inline RaceStepFn<T> race_step_fn(int mpad, int action_type, bool quiet, int phy, bool multi = false, bool dense = false)
{
    // (Command, args) actions: scripted controllers -- 13-float FULLSTATE rows, or 16-float rows of any command
    if (action_type == ADRP_ACT_MEL_FULL) return multi ? race_step_kernel<T, 0, 1, false, -1, true> : race_step_kernel<T, 0, 1, false, -1>;
    if (action_type == ADRP_ACT_MEL_CMD) return multi ? race_step_kernel<T, 0, 2, false, -1, true> : race_step_kernel<T, 0, 2, false, -1>;
    if constexpr (M<T>::fast) {
        RaceStepFn<T> f = nullptr;
        constexpr int ALL = ADRP_PHY_GND | ADRP_PHY_DRAG | ADRP_PHY_DW | ADRP_PHY_GROUND;
        if (action_type == ADRP_ACT_MEL && dense) {            // grids longer than one wave: the 128-register instances
            if (phy == ADRP_PHY_GROUND) f = race_step_special_dense<ADRP_PHY_GROUND>(mpad, quiet, multi);
            if (phy == ALL) f = race_step_special_dense<ALL>(mpad, quiet, multi);
            if (f) return f;
        }
        if (phy == ADRP_PHY_GROUND) f = multi ? race_step_special_multi<ADRP_PHY_GROUND>(mpad, quiet) : race_step_special<ADRP_PHY_GROUND>(mpad, quiet);
        if (phy == ALL) f = multi ? race_step_special_multi<ALL>(mpad, quiet) : race_step_special<ALL>(mpad, quiet);
        if (f) return f;
    }
    if (multi) return race_step_kernel<T, 0, 0, false, -1, true>;
    if constexpr (M<T>::fast) {
        switch (mpad) {
            case 1: return quiet ? race_step_kernel<T, 1, 0, true, -1> : race_step_kernel<T, 1, 0, false, -1>;
            case 2: return quiet ? race_step_kernel<T, 2, 0, true, -1> : race_step_kernel<T, 2, 0, false, -1>;
            case 4: return quiet ? race_step_kernel<T, 4, 0, true, -1> : race_step_kernel<T, 4, 0, false, -1>;
            case 8: return quiet ? race_step_kernel<T, 8, 0, true, -1> : race_step_kernel<T, 8, 0, false, -1>;
            default: break;
        }
    }
    return race_step_kernel<T, 0, 0, false, -1>;
}

template <typename T>
cudaError_t Launch<T>::race_prepare(const RaceParams<T>& RP, const KParams<T>& P)
{
    const size_t bytes = race_smem_bytes(P, RP);
    if (bytes <= 48 * 1024) return cudaSuccess;
    cudaError_t e = cudaSuccess;
    for (int v = 0; v < 8 && e == cudaSuccess; ++v)
        e = cudaFuncSetAttribute(race_step_fn<T>(P.Mpad, P.action_type, (v & 1) != 0, P.physics, (v & 2) != 0, (v & 4) != 0), cudaFuncAttributeMaxDynamicSharedMemorySize,
                                 (int)(race_smem_bytes(P, RP, kRaceBlock) > kPadBudget ? race_smem_bytes(P, RP, kRaceBlock) : kPadBudget));   // (never below launch_pdl's padding)
    if (e != cudaSuccess) return e;
    return cudaFuncSetAttribute(race_reset_kernel<T>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)bytes);
}

template <typename T>
cudaError_t Launch<T>::race_reset(const KParams<T>& P, const RaceParams<T>& RP, const KBuffers<T>& b, const RaceBuffers<T>& rb,
                                  const uint8_t* mask, const T* draws, T* obs, cudaStream_t s)
{
    race_reset_kernel<T><<<grid_for(P.Dpad), kBlock, race_smem_bytes(P, RP), s>>>(P, RP, b, rb, mask, draws, obs);
    return cudaGetLastError();
}

template <typename T>
cudaError_t Launch<T>::race_step(const KParams<T>& P, const RaceParams<T>& RP, const KBuffers<T>& b, const RaceBuffers<T>& rb,
                                 const T* action, const T* noise, T* obs, T* reward, uint8_t* te, uint8_t* tr,
                                 uint8_t* fl, cudaStream_t s)
{
    // Which register budget an instance family runs at is a property of the CONFIGURATION, never of the batch size: two
    // instances of one kernel differ at rounding level (ptxas fuses multiply-adds differently under another budget), and a
    // shard of a batch must reproduce the rows of the whole batch bit for bit (tests/test_cuda_fullsize.py).  Launches without
    // disturbance draws (QUIET) run the dense 128-register instances (C5M +5.6 %); with draws the Philox / Box-Muller state
    // makes the cap a loss (C3 -2 %).  ADRP_RACE_DENSE=0 / 1 forces one or the other (diagnostic).
    const long long threads = (long long)(P.env_end - P.env_begin) * P.Mpad;
    static const int force_dense = [] { const char* e = getenv("ADRP_RACE_DENSE"); return e ? (e[0] == '0' ? 0 : 1) : -1; }();
    const bool quiet_launch = !noise && !RP.disturbances;
    const bool dense = force_dense >= 0 ? force_dense == 1 : quiet_launch;
    launch_pdl(race_step_fn<T>(P.Mpad, P.action_type, quiet_launch, noise ? -1 : P.physics, P.K > 1 || P.act_rng, dense),
               threads, kRaceBlock,
               race_smem_bytes(P, RP, kRaceBlock), s, !(P.K > 1 || P.act_rng), P, RP, b, rb, action, noise, obs, reward, te, tr, fl);
    return cudaGetLastError();
}

template <typename T>
cudaError_t Launch<T>::dslpid_rows(T pid_gravity, T pid_kf, int model, int n, T dt, const T* pos, const T* quat, const T* vel,
                                   const T* tpos, const T* trpy, const T* tvel, const T* trates, T* ctrl, T* rpm, T* pos_e,
                                   T* yaw_e, cudaStream_t s)
{
    dslpid_rows_kernel<T><<<(unsigned)((n + 127) / 128), 128, 0, s>>>(pid_gravity, pid_kf, model, n, dt, pos, quat, vel, tpos, trpy,
                                                                      tvel, trates, ctrl, rpm, pos_e, yaw_e);
    return cudaGetLastError();
}

}  // namespace adrp
