// adrp_race.cuh -- the race env's fused control-step kernel.
//
// Replaces MultiRaceAviary.step (envs/MultiRaceAviary.py:171-270): full-state setpoint per drone, then S
// substeps of {force models + integrator (_apply_physics, :510-548), wind and motor-noise disturbances
// (:223-228,:532-544), Mellinger controller tick (control/MellingerControl.py:154-262)}, then
// _gate_progress (:471-506), _computeObs (:566-661), _collision (:552-562), _computeTerminated /
// _computeTruncated (:674-709) -- one launch, one thread per drone, the env's drones in adjacent lanes of
// one warp, the block's track instances staged in shared memory.
#pragma once
#include "adrp_mellinger.cuh"
#include "adrp_planner.cuh"
#include "adrp_step.cuh"
#include "adrp_track.cuh"

namespace adrp {

// ---------------------------------------------------------------- per-drone race state in HBM (SoA)
template <typename T>
__device__ __forceinline__ void load_mel(const RaceBuffers<T>& rb, long long D, long long t, Mel<T>& m)
{
    m.prev_rpy = {ld(rb.mel_t, D, MT_PRPY0, t), ld(rb.mel_t, D, MT_PRPY1, t), ld(rb.mel_t, D, MT_PRPY2, t)};
    m.prev_vel = {ld(rb.mel_t, D, MT_PVEL0, t), ld(rb.mel_t, D, MT_PVEL1, t), ld(rb.mel_t, D, MT_PVEL2, t)};
#pragma unroll
    for (int k = 0; k < 4; ++k) m.rpms[k] = ld(rb.mel_t, D, MT_RPM0 + k, t);
    const float* f = rb.mel_f;
#pragma unroll
    for (int k = 0; k < 3; ++k) { m.gd1[k] = ld(f, D, MF_GD1X + k, t); m.gd2[k] = ld(f, D, MF_GD2X + k, t); }
    m.fw.i_error_x = ld(f, D, MF_IEX, t); m.fw.i_error_y = ld(f, D, MF_IEY, t); m.fw.i_error_z = ld(f, D, MF_IEZ, t);
    m.fw.i_error_m_x = ld(f, D, MF_IMX, t); m.fw.i_error_m_y = ld(f, D, MF_IMY, t); m.fw.i_error_m_z = ld(f, D, MF_IMZ, t);
    m.fw.prev_omega_roll = ld(f, D, MF_POR, t); m.fw.prev_omega_pitch = ld(f, D, MF_POP, t);
    m.fw.prev_setpoint_omega_roll = ld(f, D, MF_PSR, t); m.fw.prev_setpoint_omega_pitch = ld(f, D, MF_PSP, t);
    m.fw.thrust = ld(f, D, MF_THRUST, t); m.acc_z = ld(f, D, MF_ACCZ, t);
    const int32_t* q = rb.mel_i;
    m.tick = ld(q, D, MI_TICK, t); m.tumble = ld(q, D, MI_TUMBLE, t); m.last_att = ld(q, D, MI_LAST_ATT, t);
    const int rp = ld(q, D, MI_CTL_RP, t), ye = ld(q, D, MI_CTL_Y_ERR, t);
    m.fw.roll = (float)(int)(short)(rp & 0xffff); m.fw.pitch = (float)(rp >> 16);
    m.fw.yaw = (float)(int)(short)(ye & 0xffff); m.error = ye >> 16;
}

template <typename T>
__device__ __forceinline__ void store_mel(const RaceBuffers<T>& rb, long long D, long long t, const Mel<T>& m)
{
    st(rb.mel_t, D, MT_PRPY0, t, m.prev_rpy.x); st(rb.mel_t, D, MT_PRPY1, t, m.prev_rpy.y); st(rb.mel_t, D, MT_PRPY2, t, m.prev_rpy.z);
    st(rb.mel_t, D, MT_PVEL0, t, m.prev_vel.x); st(rb.mel_t, D, MT_PVEL1, t, m.prev_vel.y); st(rb.mel_t, D, MT_PVEL2, t, m.prev_vel.z);
#pragma unroll
    for (int k = 0; k < 4; ++k) st(rb.mel_t, D, MT_RPM0 + k, t, m.rpms[k]);
    float* f = rb.mel_f;
#pragma unroll
    for (int k = 0; k < 3; ++k) { st(f, D, MF_GD1X + k, t, m.gd1[k]); st(f, D, MF_GD2X + k, t, m.gd2[k]); }
    st(f, D, MF_IEX, t, m.fw.i_error_x); st(f, D, MF_IEY, t, m.fw.i_error_y); st(f, D, MF_IEZ, t, m.fw.i_error_z);
    st(f, D, MF_IMX, t, m.fw.i_error_m_x); st(f, D, MF_IMY, t, m.fw.i_error_m_y); st(f, D, MF_IMZ, t, m.fw.i_error_m_z);
    st(f, D, MF_POR, t, m.fw.prev_omega_roll); st(f, D, MF_POP, t, m.fw.prev_omega_pitch);
    st(f, D, MF_PSR, t, m.fw.prev_setpoint_omega_roll); st(f, D, MF_PSP, t, m.fw.prev_setpoint_omega_pitch);
    st(f, D, MF_THRUST, t, m.fw.thrust); st(f, D, MF_ACCZ, t, m.acc_z);
    int32_t* q = rb.mel_i;
    st(q, D, MI_TICK, t, (int32_t)m.tick); st(q, D, MI_TUMBLE, t, (int32_t)m.tumble); st(q, D, MI_LAST_ATT, t, (int32_t)m.last_att);
    st(q, D, MI_CTL_RP, t, (int32_t)(((int)m.fw.roll & 0xffff) | ((int)m.fw.pitch << 16)));
    st(q, D, MI_CTL_Y_ERR, t, (int32_t)(((int)m.fw.yaw & 0xffff) | (m.error << 16)));
}

// ---------------------------------------------------------------- reset of one drone of one env
// MultiRaceAviary.reset (:127-167): BaseAviary.reset, track randomisation (_addObstacles, :347-403),
// _drone_init (:407-467), MellingerControl.reset (:99-150).  Uniform draws: injected (`draws`, layout
// [gates 3G | obstacles 2NO | per drone: inertia 4, pos 3, rot 3]) or Philox(seed; global env, 0, k/4, episode).
template <typename T>
__device__ __noinline__ void race_reset_drone(const KParams<T>& P, const RaceParams<T>& RP, const RaceBuffers<T>& rb,
                                              long long env, int i, int episode, const T* __restrict__ draws,
                                              Drone<T>& d, Mel<T>& m, T body[4])
{
    auto u = [&](int k) -> T {
        if (draws) return draws[env * RP.R + k];
        uint32_t o[4];
        philox4x32((uint32_t)(RP.env_offset + (unsigned long long)env), 0u, (uint32_t)(k >> 2), (uint32_t)episode, RP.seed_lo, RP.seed_hi, o);
        return u01<T>(o[k & 3]);
    };
    auto map = [](T v, T lo, T hi) -> T { return lo + v * (hi - lo); };
    if (i == 0) {                                                   // the env's track instance (:359-380)
        for (int g = 0; g < RP.G; ++g) {
            T x = RP.gate_nom[g][0], y = RP.gate_nom[g][1], yaw = RP.gate_nom[g][3];
            if (RP.rnd_track) {
                x += map(u(3 * g), RP.gate_lo, RP.gate_hi);
                y += map(u(3 * g + 1), RP.gate_lo, RP.gate_hi);
                yaw += map(u(3 * g + 2), RP.gate_lo, RP.gate_hi);
            }
            T* tr = rb.track + (long long)(kGateFields * g) * P.N + env;
            T s, c;
            M<T>::sincos(yaw, &s, &c);
            tr[0] = x; tr[P.N] = y; tr[2LL * P.N] = RP.gate_nom[g][2]; tr[3LL * P.N] = yaw; tr[4LL * P.N] = c; tr[5LL * P.N] = s;
        }
        for (int o = 0; o < RP.NO; ++o) {
            T x = RP.obst_nom[o][0], y = RP.obst_nom[o][1];
            if (RP.rnd_track) {
                x += map(u(3 * RP.G + 2 * o), RP.obst_lo, RP.obst_hi);
                y += map(u(3 * RP.G + 2 * o + 1), RP.obst_lo, RP.obst_hi);
            }
            T* tr = rb.track + (long long)(kGateFields * RP.G + kObstFields * o) * P.N + env;
            tr[0] = x; tr[P.N] = y; tr[2LL * P.N] = RP.obst_nom[o][2];
        }
    }
    // _drone_init: the property dict is mutated in the loop, so the offsets ACCUMULATE over drones (:419-424)
    const int base = 3 * RP.G + 2 * RP.NO;
    body[0] = P.pyb_mass; body[1] = P.pyb_jx; body[2] = P.pyb_jy; body[3] = P.pyb_jz;
    if (RP.rnd_inertia)
        for (int j = 0; j <= i; ++j)
            for (int k = 0; k < 4; ++k)
                body[k] = clampT(body[k] + map(u(base + 10 * j + k), RP.inertia_lo[k], RP.inertia_hi[k]), T(0), T(100));
    init_drone(P, i, d);
    if (RP.rnd_state) {                                             // :434-461
        const int b = base + 10 * i + 4;
        d.p.x += map(u(b), RP.pos_lo[0], RP.pos_hi[0]);
        d.p.y += map(u(b + 1), RP.pos_lo[1], RP.pos_hi[1]);
        d.p.z += map(u(b + 2), RP.pos_lo[2], RP.pos_hi[2]);
        const V3<T> e = {P.init_rpy[i][0] + map(u(b + 3), RP.rot_lo[0], RP.rot_hi[0]),
                         P.init_rpy[i][1] + map(u(b + 4), RP.rot_lo[1], RP.rot_hi[1]),
                         P.init_rpy[i][2] + map(u(b + 5), RP.rot_lo[2], RP.rot_hi[2])};
        d.q = canon_quat(quat_from_rpy(e));
    }
    // MellingerControl.reset(initial_obs): initial_obs predates _drone_init (:143-149)
    m.prev_rpy = {RP.mel_prev_rpy[i][0], RP.mel_prev_rpy[i][1], RP.mel_prev_rpy[i][2]};
    m.prev_vel = {T(0), T(0), T(0)};
#pragma unroll
    for (int k = 0; k < 4; ++k) m.rpms[k] = T(0);
#pragma unroll
    for (int k = 0; k < 3; ++k) m.gd1[k] = m.gd2[k] = 0.0f;
    m.fw = MelFw{};
    m.acc_z = 1.0f;                                                 // _update_state(0, pos, vel, VEC3_UP, rpy)
    m.tick = m.tumble = m.last_att = m.error = 0;
}

static_assert(kHlWords <= kHlWordsMax, "the host allocates kHlWordsMax words per padded drone");
template <typename T>
__device__ inline void load_hl(const RaceBuffers<T>& rb, long long D, long long t, HlState& h)
{
    uint32_t* w = reinterpret_cast<uint32_t*>(&h);
    for (int k = 0; k < kHlWords; ++k) w[k] = rb.hl[(long long)k * D + t];
}
template <typename T>
__device__ inline void store_hl(const RaceBuffers<T>& rb, long long D, long long t, const HlState& h)
{
    const uint32_t* w = reinterpret_cast<const uint32_t*>(&h);
    for (int k = 0; k < kHlWords; ++k) rb.hl[(long long)k * D + t] = w[k];
}

template <typename T>
__device__ __forceinline__ void store_body(const RaceBuffers<T>& rb, long long D, long long t, const T body[4])
{
#pragma unroll
    for (int k = 0; k < 4; ++k) st(rb.body, D, k, t, body[k]);
}

// MultiRaceAviary._computeObs (:566-661) for one drone.  trk: shared-memory track of the block, element
// (field f, local env e) at trk[f * EPB + e].
template <typename T>
__device__ __forceinline__ void race_write_obs(const KParams<T>& P, const RaceParams<T>& RP, const T* trk, int EPB, int e,
                                               const Drone<T>& d, const V3<T>& rpy, T hz, int current_gate, int i, int lane_base,
                                               bool valid, T* __restrict__ o, T* min_dist = nullptr)
{
    // min_dist: smallest drone-to-gate / drone-to-obstacle distance of this drone, the same gate_distance / obstacle_distance
    // values the visibility flags use -- the step kernel's contact test (_collision, :552-562) compares it with its margin
    T dmin = T(1e30);
    // distances reported for culled bodies: above both thresholds, so every comparison made with them is the exact one's
    const T thr = (RP.visibility > RP.contact_margin ? RP.visibility : RP.contact_margin) + T(0.01);
    const T far_gate = T(0.2513) + P.coll_r + thr, far_obst = T(0.1061) + P.coll_r + thr;
    const T far_gate2 = far_gate * far_gate, far_obst2 = far_obst * far_obst;
    if (valid) {
        o[0] = d.p.x; o[1] = d.p.y; o[2] = d.p.z; o[3] = rpy.x; o[4] = rpy.y; o[5] = rpy.z;
        o[6] = d.v.x; o[7] = d.v.y; o[8] = d.v.z; o[9] = d.w.x; o[10] = d.w.y; o[11] = d.w.z;
        int k = 12;
        for (int g = 0; g < RP.G; ++g) {
            const T* tg = trk + (kGateFields * g) * EPB + e;
            const T x = tg[0], y = tg[EPB], z = tg[2 * EPB], yaw = tg[3 * EPB], c = tg[4 * EPB], s = tg[5 * EPB];
            // every part of a gate lies within 0.2513 m (horizontally) of its centre: a drone further away than that plus its own
            // radius plus the larger of the two thresholds (and a centimetre for rounding) is neither in range nor in contact,
            // whatever the exact distance is -- the five prism distances are only evaluated near the gate
            const T gdx = d.p.x - x, gdy = d.p.y - y;
            T dist = far_gate;
            if (gdx * gdx + gdy * gdy <= far_gate2) dist = gate_distance(d.p, hz, P.coll_r, x, y, z, c, s, RP.gate_type[g]);
            dmin = dist < dmin ? dist : dmin;
            const bool in_range = dist < RP.visibility;                                                                  // :614-626
            o[k + 4 * g] = in_range ? x : RP.gate_nom[g][0];
            o[k + 4 * g + 1] = in_range ? y : RP.gate_nom[g][1];
            o[k + 4 * g + 2] = in_range ? z : RP.gate_nom[g][2];
            o[k + 4 * g + 3] = in_range ? yaw : RP.gate_nom[g][3];
            o[k + 4 * RP.G + g] = in_range ? T(1) : T(0);
        }
        k += 5 * RP.G;
        for (int b = 0; b < RP.NO; ++b) {
            const T* to = trk + (kGateFields * RP.G + kObstFields * b) * EPB + e;
            const T x = to[0], y = to[EPB], z = to[2 * EPB];
            const T odx = d.p.x - x, ody = d.p.y - y;                      // obstacle parts: within 0.1061 m of the axis
            T dist = far_obst;
            if (odx * odx + ody * ody <= far_obst2) dist = obstacle_distance(d.p, hz, P.coll_r, x, y, z);
            dmin = dist < dmin ? dist : dmin;
            const bool in_range = dist < RP.visibility;                                                                  // :631-643
            o[k + 3 * b] = in_range ? x : RP.obst_nom[b][0];
            o[k + 3 * b + 1] = in_range ? y : RP.obst_nom[b][1];
            o[k + 3 * b + 2] = in_range ? z : RP.obst_nom[b][2];
            o[k + 3 * RP.NO + b] = in_range ? T(1) : T(0);
        }
        k += 4 * RP.NO;
        o[k] = (T)current_gate;                                                                                          // :651
    }
    if (min_dist) *min_dist = dmin;
    if (RP.compete) {                                                                                                   // :654-659
        int k = 12 + 5 * RP.G + 4 * RP.NO + 1;
        for (int j = 0; j < P.Mpad; ++j) {
            const int src = lane_base + j;
            const T ox = __shfl_sync(0xffffffffu, d.p.x, src), oy = __shfl_sync(0xffffffffu, d.p.y, src), oz = __shfl_sync(0xffffffffu, d.p.z, src);
            const T rx = __shfl_sync(0xffffffffu, rpy.x, src), ry = __shfl_sync(0xffffffffu, rpy.y, src), rz = __shfl_sync(0xffffffffu, rpy.z, src);
            if (valid && j < P.M && j != i) {
                o[k] = ox; o[k + 1] = oy; o[k + 2] = oz; o[k + 3] = rx; o[k + 4] = ry; o[k + 5] = rz;
                k += 6;
            }
        }
    }
}

}  // namespace adrp

#include "adrp_race_step.cuh"
