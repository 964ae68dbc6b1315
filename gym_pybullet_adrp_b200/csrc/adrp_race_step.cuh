// adrp_race_step.cuh -- kernels of the race env (included by adrp_race.cuh).
#pragma once

namespace adrp {

// Block shape of race_step_kernel.  The kernel needs ~170 registers per thread, so the register file holds 12 warps
// per SM (3 blocks of 128: without the bound ptxas takes 178 and the third block is lost); ADRP_RACE_BLOCK /
// ADRP_RACE_MINB exist to measure other operating points.
#ifndef ADRP_RACE_BLOCK
#define ADRP_RACE_BLOCK 64
#endif
#ifndef ADRP_RACE_MINB
#define ADRP_RACE_MINB 6
#endif
constexpr int kRaceBlock = ADRP_RACE_BLOCK, kRaceMinBlocks = ADRP_RACE_MINB;
#ifdef ADRP_RACE_MAXNREG
#define ADRP_RACE_BOUNDS(MINB) __maxnreg__(ADRP_RACE_MAXNREG)
#else
#define ADRP_RACE_BOUNDS(MINB) __launch_bounds__(kRaceBlock, MINB)
#endif
// MINB = 8 (128 registers, 4 warps per scheduler): the DENSE instances, for grids longer than one wave -- with the loop's cold
// paths out of line the cap costs C5M a few spills and buys a third more warps (2.17e9 -> 2.29e9 drone-steps/s); a grid that
// fits one wave (C3) gains nothing from more slots and keeps the 168-register instances (measured -2 % with the cap).
constexpr int kRaceMinBlocksDense = 8;
static_assert(kBlock % kRaceBlock == 0 && kRaceBlock % 32 == 0, "host-path chunks are cut in units of kBlock threads");

// Env.reset for the masked envs (MultiRaceAviary.reset, :127-167); writes the initial observation if asked.
template <typename T>
__global__ void __launch_bounds__(kBlock)
race_reset_kernel(const __grid_constant__ KParams<T> P, const __grid_constant__ RaceParams<T> RP, const KBuffers<T> b,
                  const RaceBuffers<T> rb, const uint8_t* __restrict__ mask, const T* __restrict__ draws, T* __restrict__ obs)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* trk = reinterpret_cast<T*>(smem_raw);
    const long long t = (long long)blockIdx.x * kBlock + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const int i = (int)(t & (P.Mpad - 1));
    const long long env = t / P.Mpad;
    const int EPB = kBlock / P.Mpad, e = threadIdx.x / P.Mpad;
    const bool valid = env < P.N && i < P.M;
    const bool mine = valid && (!mask || mask[env]);
    Drone<T> d;
    Mel<T> m;
    T body[4];
    int episode = 0;
    if (mine) episode = rb.episode[env];
    __syncwarp();
    if (mine) {
        race_reset_drone(P, RP, rb, env, i, episode, draws, d, m, body);
        store_drone(b, P.Dpad, t, d);
        store_mel(rb, P.Dpad, t, m);
        store_body(rb, P.Dpad, t, body);
        if (rb.hl) {                                           // ADRP_ACT_MEL_CMD: crtpCommanderHighLevelInit + TellState
            HlState hl;
            hl_reset(P, RP, i, hl);
            store_hl(rb, P.Dpad, t, hl);
        }
        rb.gate_latch[t] = 0;
        if (i == 0) { rb.episode[env] = episode + 1; b.step_counter[env] = 0; rb.shape_gate[env] = -1; }
    }
    if (!obs) return;
    __syncthreads();                                           // track written above is read back below
    const long long env0 = (long long)blockIdx.x * EPB;
    for (int idx = threadIdx.x; idx < RP.TF * EPB; idx += kBlock) {
        const int f = idx / EPB, el = idx - f * EPB;
        trk[idx] = env0 + el < P.N ? rb.track[(long long)f * P.N + env0 + el] : T(0);
    }
    __syncthreads();
    // MultiRaceAviary.reset (:143-144): the returned observation comes from BaseAviary.reset, BEFORE
    // _drone_init randomises the drones -- the nominal INIT_XYZS / INIT_RPYS (config rpy * pi/180, :95) at rest, seen
    // against the already randomised track
    const int id = valid ? i : 0;
    d.p = {P.init_xyz[id][0], P.init_xyz[id][1], P.init_xyz[id][2]};
    d.q = canon_quat(quat_from_rpy(V3<T>{RP.mel_prev_rpy[id][0], RP.mel_prev_rpy[id][1], RP.mel_prev_rpy[id][2]}));
    d.v = d.w = d.rt = {T(0), T(0), T(0)};
    const V3<T> rpy = euler_rpy(d.q);
    const T hz = drone_half_height(P.coll_r, P.coll_h, rotmat(d.q).m22);
    T* o = obs + (valid ? env * P.M + i : 0) * P.O;
    const bool wr = mine;
    race_write_obs(P, RP, trk, EPB, e, d, rpy, hz, 0, i, lane - i, wr, o);
}

// Disturbance draws of one substep of one drone: Philox(seed; global env, 64 + drone | 128 + drone, tick, episode) -> wind
// U(lo, hi)^3 (:223-228) and motor noise N(0, std)^4 by Box-Muller (:532-544).
template <typename T>
__device__ __forceinline__ void race_draw(const RaceParams<T>& RP, long long env, int i, int tick_, int episode, T nz[7])
{
    uint32_t o0[4], o1[4];
    const uint32_t genv = (uint32_t)(RP.env_offset + (unsigned long long)env), tick = (uint32_t)tick_;
    philox4x32(genv, (uint32_t)(64 + i), tick, (uint32_t)episode, RP.seed_lo, RP.seed_hi, o0);
    philox4x32(genv, (uint32_t)(128 + i), tick, (uint32_t)episode, RP.seed_lo, RP.seed_hi, o1);
#pragma unroll
    for (int k = 0; k < 3; ++k) nz[k] = RP.wind_lo[k] + u01<T>(o0[k]) * (RP.wind_hi[k] - RP.wind_lo[k]);
    if (M<T>::fast) {
        T g0, g1, g2, g3;
        M<T>::box_muller(u01_open<T>(o1[0]), u01<T>(o1[1]), &g0, &g1);
        M<T>::box_muller(u01_open<T>(o1[2]), u01<T>(o1[3]), &g2, &g3);
        nz[3] = RP.noise_std * g0; nz[4] = RP.noise_std * g1;
        nz[5] = RP.noise_std * g2; nz[6] = RP.noise_std * g3;
    } else {
        const T r0 = M<T>::sqrt(T(-2) * M<T>::log(u01_open<T>(o1[0]))), r1 = M<T>::sqrt(T(-2) * M<T>::log(u01_open<T>(o1[2])));
        const T a0 = T(6.283185307179586) * u01<T>(o1[1]), a1 = T(6.283185307179586) * u01<T>(o1[3]);
        T s0, c0, s1, c1;
        M<T>::sincos(a0, &s0, &c0);
        M<T>::sincos(a1, &s1, &c1);
        nz[3] = RP.noise_std * r0 * c0; nz[4] = RP.noise_std * r0 * s0;
        nz[5] = RP.noise_std * r1 * c1; nz[6] = RP.noise_std * r1 * s1;
    }
}

// MultiRaceAviary.step (:171-270).  noise: NULL or (N, M, S, 7) = [wind force xyz (N), motor thrust noise x4 (N)].
// MP: lanes per env at compile time (unrolled neighbour loops), 0 = read from the params.
// ACTM: action rows -- 0: [x, y, z, yaw] (ADRP_ACT_MEL); 1: 13-float (Command.FULLSTATE, args) rows (ADRP_ACT_MEL_FULL);
// 2: 16-float (Command, args) rows of any command, the high-level commander on the device (ADRP_ACT_MEL_CMD).
// QUIET: the launch has no disturbance draws (noise == NULL and RaceParams::disturbances == 0).
// PHY_CT >= 0: physics mask known at compile time (no per-substep mask tests: +6 % on C5M); -1: read from the params.
// MULTI: adrp_step_n instances -- P.K control steps per launch (the single-step instances compile the loop away).
template <typename T, int MP, int ACTM, bool QUIET, int PHY_CT, bool MULTI = false, int MINB = kRaceMinBlocks>
__global__ void ADRP_RACE_BOUNDS(MINB)
race_step_kernel(const __grid_constant__ KParams<T> P, const __grid_constant__ RaceParams<T> RP, const KBuffers<T> b,
                 const RaceBuffers<T> rb, const T* __restrict__ action, const T* __restrict__ noise, T* __restrict__ obs,
                 T* __restrict__ reward, uint8_t* __restrict__ terminated, uint8_t* __restrict__ truncated,
                 uint8_t* __restrict__ drone_flags)
{
    extern __shared__ __align__(16) unsigned char smem_raw[];
    T* trk = reinterpret_cast<T*>(smem_raw);
    constexpr bool CMD = ACTM == 2, FULL = ACTM != 0;         // command rows carry the 13 FULLSTATE floats after the command id
    const int mpad = MP > 0 ? MP : P.Mpad;
    const long long t = (long long)P.env_begin * mpad + (long long)blockIdx.x * blockDim.x + threadIdx.x;
    const int lane = threadIdx.x & 31;
    const int i = (int)(t & (mpad - 1));
    const long long env = t / mpad;
    const int lane_base = lane - i;
    const int nthr = (int)blockDim.x;                         // 32 or 64: launch_pdl picks the size that balances the SMs
    const int EPB = nthr / mpad, e = threadIdx.x / mpad;
    const bool valid = env < P.env_end && i < P.M;
    const long long D = P.Dpad;
    const int phy = PHY_CT >= 0 ? PHY_CT : P.physics;
    const long long od = valid ? env * P.M + i : 0;

    pdl_wait();                                               // (the track below is rewritten by the previous step's auto-resets)

    Drone<T> d;
    Mel<T> m;
    T body[4] = {P.pyb_mass, P.pyb_jx, P.pyb_jy, P.pyb_jz};
    int gl = 0, counter = 0, episode = 0;
    if (valid) {
        load_drone(b, D, t, d);
        load_mel(rb, D, t, m);
#pragma unroll
        for (int k = 0; k < 4; ++k) body[k] = ld(rb.body, D, k, t);
        gl = rb.gate_latch[t];
        counter = b.step_counter[env];
        episode = rb.episode[env];
    } else {
        init_drone(P, 0, d);
        m = Mel<T>{};
    }
    HlState hl = {};
    if (CMD && valid) load_hl(rb, D, t, hl);
    int current_gate = gl & 0xffff, latch = gl >> 16;
    // RewardWrapper state (wrapper.py:78-80) of this env, held by drone 0's thread
    int sh_gate = 0;
    T sh[6] = {T(0), T(0), T(0), T(0), T(0), T(0)};
    const bool shaper = (RP.wrappers & ADRP_WRAP_REWARD) && valid && i == 0;
    if (shaper) {
        sh_gate = rb.shape_gate[env];
        if (sh_gate >= 0) {
#pragma unroll
            for (int k = 0; k < 6; ++k) sh[k] = rb.shape_vec[(long long)k * P.N + env];
        }
    }

    // ---- K control steps per launch (adrp_step_n; K = 1 for adrp_step): rigid-body, controller and latch state stay in
    // registers from one control step to the next
    int restage = 1;
    const int K = MULTI ? P.K : 1;
    for (int ks = 0; ks < K; ++ks) {
    // ---- stage the block's track instances (coalesced over envs) in shared memory: after the state loads have been issued
    // (their latency covers this), six loads in flight per thread -- one element at a time the loop was 18 exposed L2 round
    // trips per control step for level3's 36 fields (14 % of the stall samples of the C3 capture).  First read after the
    // __syncthreads that follows the substep loop.  Later control steps of the same launch re-stage only after an in-kernel
    // reset re-randomised a track of this block (the barrier also orders the previous step's reads of the old copy).
    const int kv = ks;
    if (ks > 0) restage = __syncthreads_or(restage);
    if (restage) {
        const long long env0 = (long long)P.env_begin + (long long)blockIdx.x * EPB;
        const int total = RP.TF * EPB;
        const int epb_log2 = 31 - __clz(EPB);
        for (int base = threadIdx.x; base < total; base += 6 * nthr) {
            T v[6];
#pragma unroll
            for (int u = 0; u < 6; ++u) {
                const int idx = base + u * nthr;
                const int f = idx >> epb_log2, el = idx & (EPB - 1);              // (EPB: block size / lanes per env, powers of two)
                v[u] = (idx < total && env0 + el < P.N) ? rb.track[(long long)f * P.N + env0 + el] : T(0);
            }
#pragma unroll
            for (int u = 0; u < 6; ++u)
                if (base + u * nthr < total) trk[base + u * nthr] = v[u];
        }
    }
    restage = 0;
    const bool eliminated = (latch & ADRP_FLAG_ELIMINATED) != 0;

    // ---- ndarray action -> FULLSTATE(pos, 0, 0, yaw, 0, step_counter) (:190-202)
    MelSetpoint sp = {{0.0f, 0.0f, 0.0f}, 0.0f, 1.0f, {0.0f, 0.0f, 0.0f}, {0.0f, 0.0f, 0.0f}, {0.0f, 0.0f, 0.0f}, 1.0f, 0.0f};
    bool new_setpoint = valid && !eliminated;
    const T* a = action + (long long)kv * P.act_step + od * (CMD ? 16 : (FULL ? 13 : 4));
    if (CMD && new_setpoint) {                                // any (Command, args) tuple: only FULLSTATE carries a setpoint
        new_setpoint = hl_command<T>(hl, a);
        a += 1;
    }
    if (new_setpoint) {
        T a_yaw;
        if (MULTI && !FULL && !CMD && P.act_rng) {            // a random policy's draw, made here (KParams::act_rng)
            uint32_t o[4];
            philox4x32((uint32_t)(RP.env_offset + (unsigned long long)env), (uint32_t)(192 + i), P.rng_step0 + (uint32_t)ks, 0u, RP.seed_lo, RP.seed_hi, o);
            sp.pos[0] = u11<float>(o[0]); sp.pos[1] = u11<float>(o[1]); sp.pos[2] = u11<float>(o[2]); a_yaw = u11<T>(o[3]);
        } else if (M<T>::fast && !FULL && P.act_vec) {        // one 16-byte load per drone instead of four strided ones
            const float4 a4 = *reinterpret_cast<const float4*>(a);
            sp.pos[0] = a4.x; sp.pos[1] = a4.y; sp.pos[2] = a4.z; a_yaw = (T)a4.w;
        } else {
            sp.pos[0] = (float)a[0]; sp.pos[1] = (float)a[1]; sp.pos[2] = (float)a[2]; a_yaw = a[FULL ? 9 : 3];
        }
        if (FULL) {                                           // _sendFullStateCmd (MellingerControl.py:510-527)
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                sp.vel[k] = (float)a[3 + k]; sp.acc[k] = (float)a[6 + k];
                sp.rate_deg[k] = (float)(a[10 + k] * (T)(180.0 / 3.14159265358979323846));
            }
        }
        T sy, cy;
        const T yaw = (RP.wrappers & ADRP_WRAP_OBS) ? T(0) : a_yaw;          // DroneObservationWrapper.step (wrapper.py:52-57)
        M<T>::sincos(yaw / T(2), &sy, &cy);                   // utils.get_quaternion_from_euler(0, 0, yaw)
        sp.qz = (float)sy; sp.qw = (float)cy;
        mel_setpoint_heading<T>(sp);
        if (CMD) hl.sp = sp;                                  // (yaw_abs = 0: mode.quat = modeAbs, mode.yaw = modeDisable, :536-540)
    }

    // ---- RewardWrapper.reset (wrapper.py:96-101) reads drone 0's row of the reset observation: current gate id,
    // observed pose of gate 0, position.  Evaluated on the state the episode starts from, by drone 0's thread.
    if (shaper) {
        if (sh_gate < 0) {
            sh_gate = current_gate;
            if (RP.G > 0) {
                const T* tg = rb.track + env;                  // gate 0 of this env, straight from HBM (6 reads per episode)
                const long long N = P.N;
                const T hz0 = drone_half_height(P.coll_r, P.coll_h, rotmat(d.q).m22);
                const bool in_range = gate_distance(d.p, hz0, P.coll_r, tg[0], tg[N], tg[2 * N], tg[4 * N], tg[5 * N], RP.gate_type[0]) < RP.visibility;
                sh[0] = in_range ? tg[0] : RP.gate_nom[0][0];
                sh[1] = in_range ? tg[N] : RP.gate_nom[0][1];
                sh[2] = in_range ? tg[2 * N] : RP.gate_nom[0][2];
            }
            sh[3] = d.p.x; sh[4] = d.p.y; sh[5] = d.p.z;
        }
    }
    // ---- S fused substeps (:210-254)
    // fast build: one rotation matrix per substep, computed right after the pose update and shared by the controller
    // (Euler angles, firmware attitude error) and the next substep's force models; the controller's schedule bits of
    // the whole control step fetched once
    if (M<T>::fast) d.q = unit_quat_loop(d.q);                // unit pose inside the loop (rotmat_unit, adrp_math.cuh)
    Mat3<T> R = M<T>::fast ? rotmat_unit(d.q) : rotmat_loop(d.q);
    const MelWindow win = M<T>::fast ? mel_schedule_window(RP, rb.lut, m.tick, P.S) : MelWindow{0u, false};
    // product build, Bullet free body: the angular velocity stays in the base frame for the whole control step
    const bool body_rates = M<T>::fast && !(phy & ADRP_PHY_DYN);
    V3<T> wl = {T(0), T(0), T(0)};
    if (body_rates) wl = pyb_body_rates(R, d.w);
    // the product build's instances of one physics mask with disturbances always draw them (launches with INJECTED draws run
    // the generic instance, Launch::race_step): no branch around the generator, 36 instructions and 32 B of spills less in the
    // loop (C3 7.99e8 -> 8.33e8).  Drawing one substep ahead, so that nothing waits for the generator, measured no further gain.
    constexpr bool kDrawOnly = M<T>::fast && PHY_CT >= 0 && !QUIET;
    for (int s = 0; s < P.S; ++s) {
        // disturbance draws of this substep: injected, or Philox(seed; env, drone, tick, episode)
        T nz[7] = {T(0), T(0), T(0), T(0), T(0), T(0), T(0)};
        if (QUIET) {
        } else if (!kDrawOnly && noise) {
            if (valid) {
                const T* np_ = noise + (long long)kv * P.noise_step + ((od * P.S) + s) * 7;
#pragma unroll
                for (int k = 0; k < 7; ++k) nz[k] = np_[k];
            }
        } else if (kDrawOnly || RP.disturbances) {
            race_draw<T>(RP, env, i, counter + s, episode, nz);
        }
        // _apply_physics(self.rpms, self.prev_rpms) (:510-548): forces from the substep-start state
        if (!M<T>::fast) R = rotmat_loop(d.q);
        T f[4], r2[4];
#pragma unroll
        for (int k = 0; k < 4; ++k) { r2[k] = m.rpms[k] * m.rpms[k]; f[k] = mul_rn(r2[k], P.kf); }
        if (phy & ADRP_PHY_GND) {
            if (M<T>::fast) {
                ground_effect_total(P, d, R, f, f);
            } else {
                T gnd[4];
                ground_effect(P, d, R, f, gnd);
#pragma unroll
                for (int k = 0; k < 4; ++k) f[k] += gnd[k];
            }
        }
        V3<T> fext = {T(0), T(0), T(0)};
        if (phy & ADRP_PHY_DRAG) fext = drag_force(P, d);      // prev_rpms = state row 16:20
        T dw = T(0);
        if (phy & ADRP_PHY_DW) dw = downwash<T, MP>(P, d.p, lane_base, valid);
        if (phy & ADRP_PHY_DYN) {
            T tq[4];
#pragma unroll
            for (int k = 0; k < 4; ++k) tq[k] = mul_rn(r2[k], P.km);
            dyn_substep(P, d, R, f, tq, dw, fext, s == P.S - 1);
        } else {
            // yaw torque +t0 - t1 + t2 - t3 (BaseAviary.py:703): the product build multiplies the signed sum of squares once
            // (equal commands still cancel exactly); the parity build keeps the reference's four products
            const T tz = M<T>::fast ? mul_rn((r2[0] - r2[1]) + (r2[2] - r2[3]), P.km)
                                    : mul_rn(r2[0], P.km) - mul_rn(r2[1], P.km) + mul_rn(r2[2], P.km) - mul_rn(r2[3], P.km);
            const T fz = dw + f[0] + f[1] + f[2] + f[3];
            const V3<T> tb = {mul_rn(P.prop_y[0], f[0]) + mul_rn(P.prop_y[1], f[1]) + mul_rn(P.prop_y[2], f[2]) + mul_rn(P.prop_y[3], f[3]),
                              mul_rn(-P.prop_x[0], f[0]) + mul_rn(-P.prop_x[1], f[1]) + mul_rn(-P.prop_x[2], f[2]) + mul_rn(-P.prop_x[3], f[3]),
                              tz};
            const V3<T> fw = QUIET ? fext : V3<T>{fext.x + nz[0], fext.y + nz[1], fext.z + nz[2]};   // drag + wind (:532-544)
            pyb_substep(P, phy, d, wl, R, fz, tb, fw, body[0], body[1], body[2], body[3]);
        }
        if (M<T>::fast) R = rotmat_unit(d.q);
        // controller round trip (:231-254)
        if (eliminated) {
#pragma unroll
            for (int k = 0; k < 4; ++k) { d.last_rpm[k] = T(0); m.rpms[k] = T(0); }      // :233-235
        } else {
#pragma unroll
            for (int k = 0; k < 4; ++k) d.last_rpm[k] = m.rpms[k];                       // prev_rpms[i] = rpms[i] (:251)
            mellinger_tick<T, CMD>(RP, rb.lut, m, sp, d.p, d.q, d.v, QUIET ? (const T*)nullptr : nz + 3, R, win, s, CMD ? &hl : nullptr);   // rpms[i] = recv() (:254)
        }
    }
    if (ks == K - 1) pdl_launch_dependents();             // the next launch's blocks may be scheduled during the epilogue
    if (body_rates) d.w = pyb_world_rates(R, wl);           // R: the pose after the last substep
    if (M<T>::fast) d.q = canon_quat_loop(d.q);             // once per control step in the product build (adrp_physics.cuh)
    __syncthreads();                                                                     // shared track is complete

    // ---- _gate_progress (:471-506): 7 vertical rays through the drone's current gate; the gate counts as
    // passed when the first body hit along any ray is this drone
    const Mat3<T> Rn = rotmat(d.q);
    const V3<T> axis = {Rn.m02, Rn.m12, Rn.m22};
    {
        const bool active = valid && RP.G > 0 && current_gate < RP.G;
        const int g = active ? current_gate : 0;
        const T* tg = trk + (kGateFields * g) * EPB + e;
        const T gx = RP.G > 0 ? tg[0] : T(0), gy = RP.G > 0 ? tg[EPB] : T(0);
        const T gc = RP.G > 0 ? tg[4 * EPB] : T(1), gs = RP.G > 0 ? tg[5 * EPB] : T(0);
        const T height = RP.gate_type[g] == 0 ? T(1.0) : T(0.525);                       // :479-482
        const T half_length = T(0.1875);
        const T delta_x = T(0.05) * gc, delta_y = T(0.05) * gs;
        const V3<T> dir = {T(0), T(0), (height + half_length) - (height - half_length)};
        const T hh_ = T(0.5) * P.coll_h;
        const T ray_reach = T(0.16) + M<T>::sqrt(P.coll_r * P.coll_r + hh_ * hh_), ray_reach2 = ray_reach * ray_reach;
        T best[7];
        int best_id[7];
#pragma unroll
        for (int k = 0; k < 7; ++k) { best[k] = T(-1); best_id[k] = -1; }
        for (int j = 0; j < mpad; ++j) {
            const int src = lane_base + j;
            const V3<T> cj = {__shfl_sync(0xffffffffu, d.p.x, src), __shfl_sync(0xffffffffu, d.p.y, src), __shfl_sync(0xffffffffu, d.p.z, src)};
            const V3<T> aj = {__shfl_sync(0xffffffffu, axis.x, src), __shfl_sync(0xffffffffu, axis.y, src), __shfl_sync(0xffffffffu, axis.z, src)};
            if (!active || j >= P.M) continue;
            // the seven rays are vertical and start within 0.15 m of the gate centre: a cylinder whose centre is further away
            // (horizontally) than that plus its bounding radius cannot be hit by any of them
            const T jdx = cj.x - gx, jdy = cj.y - gy;
            if (jdx * jdx + jdy * jdy > ray_reach2) continue;
#pragma unroll
            for (int k = 0; k < 7; ++k) {
                // ray order of the reference: centre, +1, -1, +2, -2, +3, -3
                const int mult = k == 0 ? 0 : ((k + 1) / 2) * ((k & 1) ? 1 : -1);
                const V3<T> fr = {gx + (T)mult * delta_x, gy + (T)mult * delta_y, height - half_length};
                const T th = ray_cylinder(fr, dir, cj, aj, P.coll_r, T(0.5) * P.coll_h);
                if (th >= T(0) && (best[k] < T(0) || th < best[k])) { best[k] = th; best_id[k] = j; }
            }
        }
        bool passed = false;
#pragma unroll
        for (int k = 0; k < 7; ++k) passed = passed || (best_id[k] == i && best[k] < T(0.9999));   // :502
        if (valid && current_gate >= RP.G) latch |= ADRP_FLAG_FINISHED;                  // :505-506 (one step late)
        if (active && passed) current_gate += 1;
    }

    // ---- observation (:566-661)
    const V3<T> rpy = euler_rpy(d.q);
    const T hz = drone_half_height(P.coll_r, P.coll_h, Rn.m22);
    T track_gap;                                             // closest gate / obstacle, from the observation's visibility tests
    T* const obs_row = obs + (long long)kv * P.obs_step + od * P.O;
    race_write_obs(P, RP, trk, EPB, e, d, rpy, hz, current_gate, i, lane_base, valid, obs_row, &track_gap);

    // ---- _computeTerminated (:674-698) with _collision (:552-562)
    int fl = 0;
    if (M<T>::abs(d.p.x) > RP.bounds_hi[0] || M<T>::abs(d.p.y) > RP.bounds_hi[1] || M<T>::abs(d.p.z) > RP.bounds_hi[2]) fl |= ADRP_FLAG_OUT_OF_BOUNDS;   // :684
    if (M<T>::abs(d.w.x) > T(20) || M<T>::abs(d.w.y) > T(20) || M<T>::abs(d.w.z) > T(20)) fl |= ADRP_FLAG_UNSTABLE;                                      // :685
    if (!(phy & ADRP_PHY_DYN)) {                               // getContactPoints reflects the last stepSimulation: never in DYN
        // any gate or obstacle closer than the margin: min over the distances <=> OR over the comparisons (same values)
        bool crashed = track_gap < RP.contact_margin;
        crashed = crashed || (d.p.z - hz < RP.contact_margin);                           // plane
        if (RP.compete) {                                                                // drones collide with each other (:153-154)
            for (int j = 0; j < mpad; ++j) {
                const int src = lane_base + j;
                const T ox = __shfl_sync(0xffffffffu, d.p.x, src), oy = __shfl_sync(0xffffffffu, d.p.y, src);
                const T oz = __shfl_sync(0xffffffffu, d.p.z, src), ohz = __shfl_sync(0xffffffffu, hz, src);
                if (j < P.M && j != i) {
                    const T dxy = gap_circle_circle(d.p.x, d.p.y, P.coll_r, ox, oy, P.coll_r);
                    crashed = crashed || combine_gaps(dxy, M<T>::abs(d.p.z - oz) - hz - ohz) < RP.contact_margin;
                }
            }
        }
        if (crashed) fl |= ADRP_FLAG_CRASHED;
    }
    if (fl) latch |= ADRP_FLAG_ELIMINATED;
    if (m.error) fl |= ADRP_FLAG_TUMBLE_STOP;
    const int not_done = valid && !(latch & (ADRP_FLAG_ELIMINATED | ADRP_FLAG_FINISHED));
    int term = group_or(not_done, mpad) == 0;
    T rew = T(0);                                                                        // :665-670
    if (shaper) {
        // RewardWrapper._compute_reward (wrapper.py:121-186) on this step's observation row of drone 0.  Declared
        // where the reference raises: info has no "task_completed" key in this fork -> drone 0's FINISHED latch;
        // gate id == number of gates has no entry in its pose table -> the target is kept.
        const T* o = obs_row;
        T r_passed = T(0), r_end = T(0);
        if (current_gate > sh_gate % 4) {
            sh_gate = current_gate;
            if (current_gate < RP.G) { sh[0] = o[12 + 4 * current_gate]; sh[1] = o[13 + 4 * current_gate]; sh[2] = o[14 + 4 * current_gate]; }
            r_passed = T(5);
        }
        if (term) r_end = (latch & ADRP_FLAG_FINISHED) ? T(10) : T(-1);
        const T dpx = sh[0] - sh[3], dpy = sh[1] - sh[4], dcx = sh[0] - d.p.x, dcy = sh[1] - d.p.y;
        const T prog_xy = M<T>::sqrt(dpx * dpx + dpy * dpy) - M<T>::sqrt(dcx * dcx + dcy * dcy);
        const T prog_z = M<T>::abs(sh[2] - sh[5]) - M<T>::abs(sh[2] - d.p.z);
        rew = prog_xy + prog_z + r_passed + r_end;
        sh[3] = d.p.x; sh[4] = d.p.y; sh[5] = d.p.z;
    }
    if (RP.wrappers & ADRP_WRAP_OBS)                                                     // wrapper.py:60-61
        term |= __shfl_sync(0xffffffffu, current_gate, lane_base) >= 2;
    const int trunc = counter >= P.trunc_steps;                                          // step_counter / PYB_FREQ > episode_len_sec (:709)
    const bool done = P.auto_reset && (term || trunc);
    if (valid && drone_flags) drone_flags[(long long)kv * P.drone_step + od] = (uint8_t)(fl | latch);
    if (valid && i == 0) {
        const long long eo = (long long)kv * P.env_step + env;
        reward[eo] = rew;
        terminated[eo] = (uint8_t)term;
        truncated[eo] = (uint8_t)trunc;
    }
    __syncwarp();
    if (valid && done) {
        // reset into copies: the out-of-line call takes addresses, and the live state must stay in registers
        Drone<T> d2;
        Mel<T> m2;
        T body2[4];
        race_reset_drone(P, RP, rb, env, i, episode, (const T*)nullptr, d2, m2, body2);
        d = d2; m = m2;
#pragma unroll
        for (int k = 0; k < 4; ++k) body[k] = body2[k];
        store_body(rb, D, t, body2);
        if (CMD) hl_reset(P, RP, i, hl);
        current_gate = 0; latch = 0;
        episode += 1;
        sh_gate = -1;
        restage = 1;                                          // this env's track was re-randomised in HBM
    }
    counter = done ? 0 : counter + P.S;                                                  // BaseAviary.py:386
    }   // control steps of this launch

    if (valid) {
        if (i == 0) { b.step_counter[env] = counter; rb.episode[env] = episode; }
        if (shaper) {
            rb.shape_gate[env] = sh_gate;
#pragma unroll
            for (int k = 0; k < 6; ++k) rb.shape_vec[(long long)k * P.N + env] = sh[k];
        }
        store_drone(b, D, t, d);
        store_mel(rb, D, t, m);
        if (CMD) store_hl(rb, D, t, hl);
        rb.gate_latch[t] = current_gate | (latch << 16);
    }
}

}  // namespace adrp
