// adrp_step.cuh -- the fused control-step kernel: one thread per drone, S substeps in registers.
//
// Replaces BaseAviary.step (envs/BaseAviary.py:262-387) for the Ctrl / Hover / MultiHover aviaries:
// action map (_preprocessAction), optional DSLPID, S x {force models, integrator}, then observation,
// reward, terminated / truncated -- one launch, state read once and written once.
#pragma once
#include "adrp_dslpid.cuh"
#include "adrp_physics.cuh"

#ifndef ADRP_PACKED_GND
#define ADRP_PACKED_GND 1
#endif
namespace adrp {

constexpr int kBlock = 128;
// Block size of step_kernel (a divisor of kBlock: the host path cuts its chunks in units of kBlock threads).
#ifndef ADRP_STEP_BLOCK
#define ADRP_STEP_BLOCK 128
#endif
#ifndef ADRP_STEP_MINB
#define ADRP_STEP_MINB 1
#endif
constexpr int kStepBlock = ADRP_STEP_BLOCK;
// Ring slots loaded before the first observation store of the RL aviaries' action-buffer copy.  The registers of the batch
// count for EVERY instance of step_kernel (the env kind is a run-time branch): 8 slots took the all-forces instances from 88 to
// 110 registers and C5 from 1.18e10 to 1.09e10 drone-steps/s.
#ifndef ADRP_RING_BATCH
#define ADRP_RING_BATCH 4
#endif
constexpr int kRingBatchGeneric = ADRP_RING_BATCH;       // (the RL-only instances, ENVK = 1, take 8)
static_assert(kBlock % kStepBlock == 0 && kStepBlock % 32 == 0, "step blocks must tile the host path's chunk unit");

template <typename T> __device__ __forceinline__ T ld(const T* base, long long D, int field, long long t) { return base[(long long)field * D + t]; }
template <typename T> __device__ __forceinline__ void st(T* base, long long D, int field, long long t, T v) { base[(long long)field * D + t] = v; }

template <typename T>
__device__ __forceinline__ void load_drone(const KBuffers<T>& b, long long D, long long t, Drone<T>& d)
{
    const T* s = b.state;
    d.p = {ld(s, D, F_PX, t), ld(s, D, F_PY, t), ld(s, D, F_PZ, t)};
    d.q = {ld(s, D, F_QX, t), ld(s, D, F_QY, t), ld(s, D, F_QZ, t), ld(s, D, F_QW, t)};
    d.v = {ld(s, D, F_VX, t), ld(s, D, F_VY, t), ld(s, D, F_VZ, t)};
    d.w = {ld(s, D, F_WX, t), ld(s, D, F_WY, t), ld(s, D, F_WZ, t)};
    d.rt = {ld(s, D, F_RX, t), ld(s, D, F_RY, t), ld(s, D, F_RZ, t)};
#pragma unroll
    for (int k = 0; k < 4; ++k) d.last_rpm[k] = ld(s, D, F_RPM0 + k, t);
}

template <typename T>
__device__ __forceinline__ void store_drone(const KBuffers<T>& b, long long D, long long t, const Drone<T>& d)
{
    T* s = b.state;
    st(s, D, F_PX, t, d.p.x); st(s, D, F_PY, t, d.p.y); st(s, D, F_PZ, t, d.p.z);
    st(s, D, F_QX, t, d.q.x); st(s, D, F_QY, t, d.q.y); st(s, D, F_QZ, t, d.q.z); st(s, D, F_QW, t, d.q.w);
    st(s, D, F_VX, t, d.v.x); st(s, D, F_VY, t, d.v.y); st(s, D, F_VZ, t, d.v.z);
    st(s, D, F_WX, t, d.w.x); st(s, D, F_WY, t, d.w.y); st(s, D, F_WZ, t, d.w.z);
    st(s, D, F_RX, t, d.rt.x); st(s, D, F_RY, t, d.rt.y); st(s, D, F_RZ, t, d.rt.z);
#pragma unroll
    for (int k = 0; k < 4; ++k) st(s, D, F_RPM0 + k, t, d.last_rpm[k]);
}

template <typename T>
__device__ __forceinline__ void init_drone(const KParams<T>& P, int i, Drone<T>& d)
{
    d.p = {P.init_xyz[i][0], P.init_xyz[i][1], P.init_xyz[i][2]};
    d.q = {P.init_quat[i][0], P.init_quat[i][1], P.init_quat[i][2], P.init_quat[i][3]};
    d.v = {P.init_vel[i][0], P.init_vel[i][1], P.init_vel[i][2]};
    d.w = {P.init_pqr[i][0], P.init_pqr[i][1], P.init_pqr[i][2]};
    d.rt = {T(0), T(0), T(0)};
#pragma unroll
    for (int k = 0; k < 4; ++k) d.last_rpm[k] = T(0);
}

// sum over the Mpad lanes of one env (Mpad is a power of two <= 32)
template <typename T> __device__ __forceinline__ T group_sum(T v, int Mpad)
{
    for (int o = Mpad >> 1; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
// Programmatic dependent launch (sm_90+): the step kernels are launched with programmatic stream serialisation, so the
// blocks of control step n+1 are scheduled while the last wave of step n drains; they run their index arithmetic and then
// wait here for step n to complete and flush (no global access before pdl_wait).  Both are no-ops for a kernel launched
// without the attribute or behind a non-kernel operation.
__device__ __forceinline__ void pdl_launch_dependents() { asm volatile("griddepcontrol.launch_dependents;"); }
__device__ __forceinline__ void pdl_wait() { asm volatile("griddepcontrol.wait;" ::: "memory"); }

// Action ring of the RL aviaries: element (slot, component k) of padded drone t.  Slot-major, then drone, then component, so
// the A components of one slot are adjacent: a 4-component action is one 16-byte access, coalesced over the warp.
__device__ __forceinline__ long long ring_idx(long long D, int A, int slot, int k, long long t)
{
    return ((long long)slot * D + t) * A + k;
}

__device__ __forceinline__ int group_or(int v, int Mpad)
{
    for (int o = Mpad >> 1; o > 0; o >>= 1) v |= __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// PHY_CT >= 0: physics mask known at compile time (specialised kernels); -1: read from the params.
// ACTC: 0 = RPM-type actions, 1 = DSLPID-type actions.  MP: lanes per env (0 = read from the params).
// ENVK: -1 = env kind read from the params; 0 = CtrlAviary-kind (20-float rows, no reward / action buffer) compiled in
// (2: the same with the 12-column row of adrp_config.obs_columns, n-step instances only);
// 1 = an RL aviary (Hover / MultiHover, told apart at run time) compiled in.  The epilogues of the two families differ
// a lot (5 stores vs. the action-buffer copy), and the registers of the larger one count for the whole kernel.
// MULTI: adrp_step_n instances -- P.K control steps per launch (the single-step instances compile the loop away and are
// the code they were before it existed: the loop costs the all-forces instances 12 registers, i.e. one resident block).
// DM: prop layout compiled in (0: X -- CF2X / RACE; -1: drone model read from the params).
// Launch bounds: the specialised single-step instances (DM >= 0) ask for 6 resident blocks of 128 (<= 85 registers): the all-forces
// Bullet instance drops from 91 to 85 without spills and C4 gains 3 %; the adrp_step_n instances need their registers for the
// control-step loop (the same bound costs them 2 - 5 %), the generic ones for the DSLPID state.
template <typename T, int PHY_CT, int ACTC, int MP, int ENVK = -1, bool MULTI = false, int DM = -1>
__global__ void __launch_bounds__(kStepBlock, (DM >= 0 && !MULTI && ADRP_STEP_MINB < 6) ? 6 : ADRP_STEP_MINB)
step_kernel(const __grid_constant__ KParams<T> P, const KBuffers<T> b, const T* __restrict__ action,
            T* __restrict__ obs, T* __restrict__ reward, uint8_t* __restrict__ terminated,
            uint8_t* __restrict__ truncated, uint8_t* __restrict__ drone_flags)
{
    const int mpad = MP > 0 ? MP : P.Mpad;
    const long long t = (long long)P.env_begin * mpad + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const int i = (int)(t & (mpad - 1));
    const long long env = t / mpad;
    const int lane_base = lane - i;
    const bool valid = env < P.env_end && i < P.M;
    const long long D = P.Dpad;
    const int phy = PHY_CT >= 0 ? PHY_CT : P.physics;
    const bool is_ctrl = (ENVK == 0 || ENVK == 2) ? true : (ENVK == 1 ? false : P.env_kind == ADRP_ENV_CTRL);
    constexpr int kRingBatch = ENVK == 1 ? 8 : kRingBatchGeneric;
    const long long od = valid ? env * P.M + i : 0;      // row of this drone in (N, M, .) tensors
    pdl_wait();

    Drone<T> d;
    if (valid) load_drone(b, D, t, d); else init_drone(P, 0, d);
    PidState<T> pc;
    if (ACTC == 1) {
        const T* pb = b.pid;
        pc.ipos = {ld(pb, D, 0, t), ld(pb, D, 1, t), ld(pb, D, 2, t)};
        pc.irpy = {ld(pb, D, 3, t), ld(pb, D, 4, t), ld(pb, D, 5, t)};
        pc.last_rpy = {ld(pb, D, 6, t), ld(pb, D, 7, t), ld(pb, D, 8, t)};
    }
    // (single-step instances fetch these two where they are first used, after the substep loop: two registers less in it)
    int counter = 0, head = 0;
    if (MULTI) {
        counter = env < P.env_end ? b.step_counter[env] : 0;
        head = (!is_ctrl && valid) ? b.ring_head[env] : 0;
    }

    // ---- K control steps per launch (adrp_step_n; K = 1 for adrp_step): the rigid-body / controller state, the step counter
    // and the ring head stay in registers from one control step to the next
    const int K = MULTI ? P.K : 1;
    for (int ks = 0; ks < K; ++ks) {
    const int kv = ks;
    // ---- action -> RPM (CtrlAviary.py:140; BaseRLAviary.py:160-239)
    T rpm[4];
    const T* a = action + (long long)kv * P.act_step + od * P.A;
    constexpr int AMAX = ACTC == 0 ? 4 : 12;             // RPM-type actions have at most 4 components
    T act_in[AMAX];
    // a random policy's draw, made here (KParams::act_rng).  Not compiled into the CtrlAviary-kind instances (their actions are
    // RPMs, not Box(-1, 1) samples; the draw's registers cost the C4 / C5 n-step instances 3 - 5 %): such a handle runs the
    // generic instance for it (Launch::step)
    if (MULTI && ENVK != 0 && ENVK != 2 && P.act_rng) {
        uint32_t o[4];
        philox4x32((uint32_t)(P.env_offset + (unsigned long long)env), (uint32_t)(192 + i), P.rng_step0 + (uint32_t)ks, 0u, P.rng_lo, P.rng_hi, o);
#pragma unroll
        for (int k = 0; k < AMAX; ++k) act_in[k] = (k < 4 && k < P.A) ? u11<T>(o[k & 3]) : T(0);
    } else if (M<T>::fast && P.act_vec && AMAX == 4) {        // one 16-byte load per drone instead of four strided ones
        const float4 a4 = valid ? *reinterpret_cast<const float4*>(a) : make_float4(0.0f, 0.0f, 0.0f, 0.0f);
        act_in[0] = (T)a4.x; act_in[1] = (T)a4.y; act_in[2] = (T)a4.z; act_in[3] = (T)a4.w;
    } else {
#pragma unroll
        for (int k = 0; k < AMAX; ++k) act_in[k] = (valid && k < P.A) ? a[k] : T(0);
    }
    if (ACTC == 0) {
        if (P.action_type == ADRP_ACT_RPM_RAW) {
#pragma unroll
            for (int k = 0; k < 4; ++k) rpm[k] = clampT(act_in[k], T(0), P.max_rpm);
        } else {
            // NumPy evaluates (1 + 0.05*a) on the float32 action before promoting (BaseRLAviary.py:192)
#pragma unroll
            for (int k = 0; k < 4; ++k) {
                const float av = (float)(P.action_type == ADRP_ACT_ONE_D_RPM ? act_in[0] : act_in[k]);
                rpm[k] = P.hover_rpm * (T)__fadd_rn(1.0f, __fmul_rn(0.05f, av));
            }
        }
    } else {
        V3<T> tpos, trpy = {T(0), T(0), T(0)}, tvel = {T(0), T(0), T(0)}, trates = {T(0), T(0), T(0)};
        if (P.action_type == ADRP_ACT_PID) {
            // BaseAviary._calculateNextStep (:1112-1154), step_size = 1
            const V3<T> dir = {act_in[0] - d.p.x, act_in[1] - d.p.y, act_in[2] - d.p.z};
            const T dist = M<T>::sqrt(dot(dir, dir));
            if (dist <= T(1)) tpos = {act_in[0], act_in[1], act_in[2]};
            else tpos = {d.p.x + dir.x / dist * T(1), d.p.y + dir.y / dist * T(1), d.p.z + dir.z / dist * T(1)};
        } else if (P.action_type == ADRP_ACT_ONE_D_PID) {
            tpos = {d.p.x, d.p.y, d.p.z + T(0.1) * act_in[0]};                   // :233
        } else if (P.action_type == ADRP_ACT_VEL) {
            const T nrm = M<T>::sqrt(act_in[0] * act_in[0] + act_in[1] * act_in[1] + act_in[2] * act_in[2]);   // :208-223
            const T sc = P.speed_limit * M<T>::abs(act_in[3]);
            tpos = d.p;
            trpy.z = euler_rpy(d.q).z;
            if (nrm != T(0)) tvel = {sc * (act_in[0] / nrm), sc * (act_in[1] / nrm), sc * (act_in[2] / nrm)};
        } else {                                                                 // ADRP_ACT_PID_FULL
            tpos = {act_in[0], act_in[1], act_in[2]}; trpy = {act_in[3], act_in[4], act_in[5]};
            tvel = {act_in[6], act_in[7], act_in[8]}; trates = {act_in[9], act_in[10], act_in[11]};
        }
        V3<T> pos_e; T yaw_e;
        dslpid_control(P.pid_gravity, P.pid_kf, P.pid_model, P.ctrl_dt, d.p, d.q, d.v, tpos, trpy, tvel, trates,
                       pc, rpm, pos_e, yaw_e);
    }

    // ---- S fused substeps (BaseAviary.py:347-376)
    T f0[4], tq[4];
#pragma unroll
    for (int k = 0; k < 4; ++k) { f0[k] = mul_rn(rpm[k] * rpm[k], P.kf); tq[k] = mul_rn(rpm[k] * rpm[k], P.km); }
    const T drag_sum_prev = T(0.10471975511965977) * ((d.last_rpm[0] + d.last_rpm[1]) + (d.last_rpm[2] + d.last_rpm[3]));   // 2 pi / 60
    const T drag_sum_cur = T(0.10471975511965977) * ((rpm[0] + rpm[1]) + (rpm[2] + rpm[3]));
    // product build, Bullet free body: the angular velocity stays in the base frame for the whole control step
    const bool body_rates = M<T>::fast && !(phy & ADRP_PHY_DYN);
    V3<T> wl = {T(0), T(0), T(0)};
    // product build: unit pose inside the loop (see rotmat_unit); explicit model: drag as a gain on the old velocity
    if (M<T>::fast) d.q = unit_quat_loop(d.q);
    if (body_rates) wl = pyb_body_rates(rotmat_unit(d.q), d.w);
    const bool dyn_fast = M<T>::fast && (phy & ADRP_PHY_DYN);
    T axy = T(1), az = T(1), axy_cur = T(1), az_cur = T(1);
    if (dyn_fast) {
        if (phy & ADRP_PHY_DRAG) {                         // last_clipped_action: the previous step's RPM during the first substep
            const T sp = (d.last_rpm[0] + d.last_rpm[1]) + (d.last_rpm[2] + d.last_rpm[3]), sc = (rpm[0] + rpm[1]) + (rpm[2] + rpm[3]);
            axy = T(1) - P.k_drag_xy * sp; az = T(1) - P.k_drag_z * sp;
            axy_cur = T(1) - P.k_drag_xy * sc; az_cur = T(1) - P.k_drag_z * sc;
        }
    }
    // the packed fp32 forms of the force models (adrp_physics.cuh): in the n-step instances and the DYN single-step ones
    // (measured per instance, profiles/r02zm_* / r02zo_*: the ground effect's packed form costs C4's n-step instance 2.6 %)
    constexpr bool kPacked = MULTI || (PHY_CT >= 0 && (PHY_CT & ADRP_PHY_DYN) != 0);
    constexpr bool kPackedGnd = PHY_CT >= 0 && (PHY_CT & ADRP_PHY_DYN) != 0;
    for (int s = 0; s < P.S; ++s) {
        const Mat3<T> R = M<T>::fast ? rotmat_unit(d.q) : rotmat_loop(d.q);
        T f[4] = {f0[0], f0[1], f0[2], f0[3]};
        if (phy & ADRP_PHY_GND) {
            if constexpr (M<T>::fast && kPackedGnd && ADRP_PACKED_GND) {
                ground_effect_total2(P, d, R, f0, f);
            } else if (M<T>::fast) {
                ground_effect_total(P, d, R, f0, f);
            } else {
                T gnd[4];
                ground_effect(P, d, R, f0, gnd);
#pragma unroll
                for (int k = 0; k < 4; ++k) f[k] += gnd[k];
            }
        }
        V3<T> fext = {T(0), T(0), T(0)};
        if ((phy & ADRP_PHY_DRAG) && !dyn_fast) {          // last_clipped_action: the previous step's RPM during the first substep (:363,:376)
            if (M<T>::fast) fext = drag_force_sum(P, d, s == 0 ? drag_sum_prev : drag_sum_cur);
            else fext = drag_force(P, d);
        }
        T dw = T(0);
        if (phy & ADRP_PHY_DW) dw = downwash<T, MP, kPacked>(P, d.p, lane_base, valid);
        if (dyn_fast) {
            dyn_substep_fast<T, DM>(P, d, R, f, tq, dw, axy, az, s == P.S - 1);
            axy = axy_cur; az = az_cur;
        } else if (phy & ADRP_PHY_DYN) {
            dyn_substep(P, d, R, f, tq, dw, fext, s == P.S - 1);
        } else {
            const T fz = (f[0] + f[1] + f[2] + f[3]) + dw;
            const V3<T> tb = {mul_rn(P.prop_y[0], f[0]) + mul_rn(P.prop_y[1], f[1]) + mul_rn(P.prop_y[2], f[2]) + mul_rn(P.prop_y[3], f[3]),
                              -(mul_rn(P.prop_x[0], f[0]) + mul_rn(P.prop_x[1], f[1]) + mul_rn(P.prop_x[2], f[2]) + mul_rn(P.prop_x[3], f[3])),
                              tq[0] - tq[1] + tq[2] - tq[3]};                   // :703
            pyb_substep(P, phy, d, wl, R, fz, tb, fext, P.pyb_mass, P.pyb_jx, P.pyb_jy, P.pyb_jz);
        }
        if (!M<T>::fast) {
#pragma unroll
            for (int k = 0; k < 4; ++k) d.last_rpm[k] = rpm[k];                  // :376
        }
    }
    if (M<T>::fast) {
#pragma unroll
        for (int k = 0; k < 4; ++k) d.last_rpm[k] = rpm[k];                      // :376 (S >= 1)
    }
    if (ks == K - 1) pdl_launch_dependents();          // the next launch's blocks may be scheduled while this one writes out
    if (M<T>::fast) d.q = canon_quat_loop(d.q);          // once per control step in the product build (adrp_physics.cuh)
    if (body_rates) d.w = pyb_world_rates(rotmat_unit(d.q), wl);     // (d.q: normalised on the line above)

    // ---- observation (CtrlAviary.py:106-117; BaseRLAviary.py:284-322).  CtrlAviary has no flag that depends on the
    // Euler angles, so the product build may use the loop's polynomial atan2 there (1 ulp); the RL aviaries threshold
    // |roll|, |pitch| and keep libm, so that their flags stay a bit-exact function of the stored state
    const V3<T> rpy = is_ctrl ? euler_rpy_loop(d.q) : euler_rpy(d.q);
    T* o = obs + (long long)kv * P.obs_step + od * P.O;
    // fp32 build: rows whose length is a multiple of 4 floats go out as 16-byte stores (a warp's scalar store of one
    // column touches 32 sectors for 128 useful bytes; the float4 form moves 4x the payload per sector access)
    const bool vec = M<T>::fast && P.obs_vec;
    // (the run-time choice of the 12-column row is not compiled into the CtrlAviary-kind n-step instances: there it moves ptxas
    // from 93 to 84 registers and C5's adrp_step_n rate from 1.468e10 to 1.432e10; ENVK = 2 are the n-step instances of that row)
    constexpr bool kObs12 = !(MULTI && ENVK == 0);
    const bool row12 = ENVK == 2 ? true : (kObs12 && P.O == 12);
    if (is_ctrl) {
        if (valid && vec) {
            float4* o4 = reinterpret_cast<float4*>(o);
            const float4 kin1 = make_float4((float)rpy.y, (float)rpy.z, (float)d.v.x, (float)d.v.y);
            const float4 kin2 = make_float4((float)d.v.z, (float)d.w.x, (float)d.w.y, (float)d.w.z);
            if (row12) {                                       // adrp_config.obs_columns = 12: the kinematic columns only
                o4[0] = make_float4((float)d.p.x, (float)d.p.y, (float)d.p.z, (float)rpy.x);
                o4[1] = kin1;
                o4[2] = kin2;
            } else {
                o4[0] = make_float4((float)d.p.x, (float)d.p.y, (float)d.p.z, (float)d.q.x);
                o4[1] = make_float4((float)d.q.y, (float)d.q.z, (float)d.q.w, (float)rpy.x);
                o4[2] = kin1;
                o4[3] = kin2;
                o4[4] = make_float4((float)d.last_rpm[0], (float)d.last_rpm[1], (float)d.last_rpm[2], (float)d.last_rpm[3]);
            }
        } else if (valid && row12) {
            o[0] = d.p.x; o[1] = d.p.y; o[2] = d.p.z; o[3] = rpy.x; o[4] = rpy.y; o[5] = rpy.z;
            o[6] = d.v.x; o[7] = d.v.y; o[8] = d.v.z; o[9] = d.w.x; o[10] = d.w.y; o[11] = d.w.z;
        } else if (valid) {
            o[0] = d.p.x; o[1] = d.p.y; o[2] = d.p.z; o[3] = d.q.x; o[4] = d.q.y; o[5] = d.q.z; o[6] = d.q.w;
            o[7] = rpy.x; o[8] = rpy.y; o[9] = rpy.z; o[10] = d.v.x; o[11] = d.v.y; o[12] = d.v.z;
            o[13] = d.w.x; o[14] = d.w.y; o[15] = d.w.z;
#pragma unroll
            for (int k = 0; k < 4; ++k) o[16 + k] = d.last_rpm[k];
        }
    } else if (valid) {
        if (vec) {
            float4* o4 = reinterpret_cast<float4*>(o);
            o4[0] = make_float4((float)d.p.x, (float)d.p.y, (float)d.p.z, (float)rpy.x);
            o4[1] = make_float4((float)rpy.y, (float)rpy.z, (float)d.v.x, (float)d.v.y);
            o4[2] = make_float4((float)d.v.z, (float)d.w.x, (float)d.w.y, (float)d.w.z);
        } else {
            o[0] = d.p.x; o[1] = d.p.y; o[2] = d.p.z; o[3] = rpy.x; o[4] = rpy.y; o[5] = rpy.z;
            o[6] = d.v.x; o[7] = d.v.y; o[8] = d.v.z; o[9] = d.w.x; o[10] = d.w.y; o[11] = d.w.z;
        }
        // action buffer (deque(maxlen=ctrl_freq//2), BaseRLAviary.py:66-67,187): the newest action
        // overwrites the oldest ring slot; the observation lists the slots oldest -> newest
        if (!MULTI) head = b.ring_head[env];
        const int nh = head + 1 == P.B ? 0 : head + 1;
        if (vec) {                                             // P.A == 4: one 16-byte access per ring slot
            float4* r4 = reinterpret_cast<float4*>(b.ring);
            float4* o4 = reinterpret_cast<float4*>(o) + 3;
            const float4 newest = make_float4((float)act_in[0], (float)act_in[1], (float)act_in[2], (float)act_in[3]);
            r4[(long long)head * D + t] = newest;
            // kRingBatch slots in flight before the first store (the ring may alias the observation as far as the compiler knows:
            // slot by slot the loop is one L2 round trip per slot, 40 % of the warp's time in the C2 capture); the newest slot
            // comes from the registers
            for (int j0 = 0; j0 < P.B; j0 += kRingBatch) {
                float4 v[kRingBatch];
#pragma unroll
                for (int u = 0; u < kRingBatch; ++u) {
                    int slot = nh + j0 + u;
                    if (slot >= P.B) slot -= P.B;
                    if (j0 + u < P.B - 1) v[u] = r4[(long long)slot * D + t];
                    else v[u] = newest;
                }
#pragma unroll
                for (int u = 0; u < kRingBatch; ++u)
                    if (j0 + u < P.B) o4[j0 + u] = v[u];
            }
        } else {
#pragma unroll
            for (int k = 0; k < AMAX; ++k)
                if (k < P.A) b.ring[ring_idx(D, P.A, head, k, t)] = act_in[k];
            for (int j = 0; j < P.B; ++j) {
                int slot = nh + j;
                if (slot >= P.B) slot -= P.B;
                for (int k = 0; k < P.A; ++k) o[12 + j * P.A + k] = b.ring[ring_idx(D, P.A, slot, k, t)];
            }
        }
        head = nh;
    }

    // ---- reward and flags (HoverAviary.py:68-117; MultiHoverAviary.py:75-130; CtrlAviary.py:144-200)
    T rew = T(-1);
    int term = 0, trunc = 0;
    if (!MULTI) counter = env < P.env_end ? b.step_counter[env] : 0;
    if (!is_ctrl) {
        T r = T(0), dist = T(0);
        int tr = 0;
        if (valid) {
            const V3<T> e = {P.target[i][0] - d.p.x, P.target[i][1] - d.p.y, P.target[i][2] - d.p.z};
            dist = M<T>::sqrt(dot(e, e));
            const T d2 = dist * dist;
            r = M<T>::fmax(T(0), T(2) - d2 * d2);
            tr = M<T>::abs(d.p.x) > P.hover_lim || M<T>::abs(d.p.y) > P.hover_lim || d.p.z > T(2) ||
                 M<T>::abs(rpy.x) > T(.4) || M<T>::abs(rpy.y) > T(.4);
        }
        rew = group_sum(r, mpad);
        term = group_sum(dist, mpad) < T(.0001);
        trunc = group_or(tr, mpad) | (counter >= P.trunc_steps);
    }
    const bool done = P.auto_reset && (term || trunc);
    if (valid && i == 0) {
        const long long eo = (long long)kv * P.env_step + env;
        reward[eo] = rew;
        terminated[eo] = (uint8_t)term;
        truncated[eo] = (uint8_t)trunc;
    }
    counter = done ? 0 : counter + P.S;                                          // BaseAviary.py:386
    if (valid && drone_flags) drone_flags[(long long)kv * P.drone_step + od] = 0;
    if (done) {                                          // the next episode starts from the initial pose (auto_reset)
        init_drone(P, valid ? i : 0, d);
        pc.ipos = pc.irpy = pc.last_rpy = {T(0), T(0), T(0)};
    }
    }   // control steps of this launch

    // ---- write the state back
    if (valid) {
        if (i == 0) {
            b.step_counter[env] = counter;
            if (!is_ctrl) b.ring_head[env] = head;         // the action buffer survives a reset (BaseRLAviary.py:66-67)
        }
        store_drone(b, D, t, d);
        if (ACTC == 1) {
            T* pb = b.pid;
            st(pb, D, 0, t, pc.ipos.x); st(pb, D, 1, t, pc.ipos.y); st(pb, D, 2, t, pc.ipos.z);
            st(pb, D, 3, t, pc.irpy.x); st(pb, D, 4, t, pc.irpy.y); st(pb, D, 5, t, pc.irpy.z);
            st(pb, D, 6, t, pc.last_rpy.x); st(pb, D, 7, t, pc.last_rpy.y); st(pb, D, 8, t, pc.last_rpy.z);
        }
    }
}

// (N, M, 20) rows <-> SoA state
template <typename T>
__global__ void set_state_kernel(const KParams<T> P, KBuffers<T> b, const T* __restrict__ rows)
{
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int i = (int)(t & (P.Mpad - 1));
    const long long env = t / P.Mpad;
    if (env >= P.N || i >= P.M) return;
    const T* r = rows + (env * P.M + i) * ADRP_STATE_DIM;
    for (int f = 0; f < F_COUNT; ++f) b.state[(long long)f * P.Dpad + t] = r[f];
}

template <typename T>
__global__ void get_state_kernel(const KParams<T> P, KBuffers<T> b, T* __restrict__ rows)
{
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int i = (int)(t & (P.Mpad - 1));
    const long long env = t / P.Mpad;
    if (env >= P.N || i >= P.M) return;
    T* r = rows + (env * P.M + i) * ADRP_STATE_DIM;
    for (int f = 0; f < F_COUNT; ++f) r[f] = b.state[(long long)f * P.Dpad + t];
}

// Env.reset (BaseAviary.py:223-258) for the masked envs; writes the initial observation if asked.
template <typename T>
__global__ void reset_kernel(const KParams<T> P, KBuffers<T> b, const uint8_t* __restrict__ mask, T* __restrict__ obs)
{
    const long long t = (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int i = (int)(t & (P.Mpad - 1));
    const long long env = t / P.Mpad;
    if (env >= P.N || i >= P.M) return;
    if (mask && !mask[env]) return;
    Drone<T> d;
    init_drone(P, i, d);
    store_drone(b, P.Dpad, t, d);
    for (int k = 0; k < 9; ++k) b.pid[(long long)k * P.Dpad + t] = T(0);
    // the action buffer is filled once in BaseRLAviary.__init__ (:66-67) and reset() never touches it: the ring (zeroed
    // by adrp_create) and its head survive, so the observation below and the next B - 1 steps still list the previous
    // episode's actions (golden hover_pyb_rpm_reset)
    if (i == 0) b.step_counter[env] = 0;
    if (obs) {
        T* o = obs + (env * P.M + i) * P.O;
        const V3<T> rpy = euler_rpy(d.q);
        if (P.env_kind == ADRP_ENV_CTRL && P.O == 12) {        // adrp_config.obs_columns = 12
            o[0] = d.p.x; o[1] = d.p.y; o[2] = d.p.z; o[3] = rpy.x; o[4] = rpy.y; o[5] = rpy.z;
            o[6] = d.v.x; o[7] = d.v.y; o[8] = d.v.z; o[9] = d.w.x; o[10] = d.w.y; o[11] = d.w.z;
        } else if (P.env_kind == ADRP_ENV_CTRL) {
            o[0] = d.p.x; o[1] = d.p.y; o[2] = d.p.z; o[3] = d.q.x; o[4] = d.q.y; o[5] = d.q.z; o[6] = d.q.w;
            o[7] = rpy.x; o[8] = rpy.y; o[9] = rpy.z; o[10] = d.v.x; o[11] = d.v.y; o[12] = d.v.z;
            o[13] = d.w.x; o[14] = d.w.y; o[15] = d.w.z; o[16] = o[17] = o[18] = o[19] = T(0);
        } else {
            o[0] = d.p.x; o[1] = d.p.y; o[2] = d.p.z; o[3] = rpy.x; o[4] = rpy.y; o[5] = rpy.z;
            o[6] = d.v.x; o[7] = d.v.y; o[8] = d.v.z; o[9] = d.w.x; o[10] = d.w.y; o[11] = d.w.z;
            const int head = b.ring_head[env];                 // oldest slot first (BaseRLAviary.py:315-322)
            for (int j = 0; j < P.B; ++j) {
                const int slot = head + j >= P.B ? head + j - P.B : head + j;
                for (int k = 0; k < P.A; ++k) o[12 + j * P.A + k] = b.ring[ring_idx(P.Dpad, P.A, slot, k, t)];
            }
        }
    }
}

}  // namespace adrp
