// adrp_params.cuh -- per-handle constants handed to the kernels by value (__grid_constant__ -> constant bank).
#pragma once
#include <stdint.h>
#include "../../include/adrp.h"

namespace adrp {

// State in HBM: structure of arrays, one array of length Dpad = N * Mpad per scalar field, in the
// order of ADRP_STATE_DIM: every warp-wide load/store of a field is one fully coalesced line.
enum Field {
    F_PX = 0, F_PY, F_PZ, F_QX, F_QY, F_QZ, F_QW, F_VX, F_VY, F_VZ, F_WX, F_WY, F_WZ,
    F_RX, F_RY, F_RZ, F_RPM0, F_RPM1, F_RPM2, F_RPM3, F_COUNT
};

template <typename T>
struct KParams {
    // geometry of the batch
    int N, M, Mpad, S, A, O, B;           // envs, drones/env, lanes/env, substeps, action dim, obs dim, action-buffer length
    int physics, drone_model, action_type, env_kind, race_mode, auto_reset, pid_model;
    long long Dpad;                       // N * Mpad
    int env_begin, env_end;               // env range of THIS launch (chunked host path); the whole shard otherwise
    int obs_vec;                          // this launch may write observation rows with 16-byte stores (fp32, rows and base 16 B aligned)
    int act_vec;                          // ... and read 4-float action rows with one 16-byte load
    // adrp_step_n: K control steps in this launch (state in registers across them); element strides from one step to the
    // next of the action / observation / per-env (reward, terminated, truncated) / per-drone (flags) / noise buffers, 0 = the
    // same buffer every step
    int K;
    long long act_step, obs_step, env_step, drone_step, noise_step;
    int pyb_freq;
    int trunc_steps;                      // smallest step_counter with step_counter/PYB_FREQ > EPISODE_LEN_SEC
    // BaseAviary constants (envs/BaseAviary.py:74-128)
    T mass, inv_mass, arm_x, kf, km, jx, jy, jz, jinvx, jinvy, jinvz, gravity, hover_rpm, max_rpm;
    T gnd_eff_coeff, prop_radius, gnd_eff_h_clip, drag_xy, drag_z, dw1, dw2, dw3, speed_limit;
    // fp32 product build, downwash in base-2 form: term = 2^(dwf_l2c - dxy^2 / (dwf_a dz + dwf_b)^2) / dz^2 with
    // dwf_l2c = log2(dw1 (prop_radius/4)^2) and (dwf_a, dwf_b) = (dw2, dw3) / sqrt(log2(e) / 2)
    T dwf_a, dwf_b, dwf_l2c, half_dt, qr4;
    // fp32 product build, explicit model (dyn_substep_fast), folded on the host in double: body-rate update
    //   rt += k_t * (signed prop-force sums) - k_g * (products of the other two rates)
    // with k_t = dt J^-1 arm (z: dt Jz^-1, sign of the RACE model folded in), k_g = dt J^-1 (J_next - J_prev);
    // velocity update v = v * (1 - k_drag * sum_rpm) + (dt / m) thrust * R[:, 2] - (0, 0, dt g)
    T k_tx, k_ty, k_tz, k_gx, k_gy, k_gz, k_m, k_grav, k_drag_xy, k_drag_z;
    // ... Bullet free body: dt * damping coefficients; plane_reach: no part of the collision cylinder of a drone whose centre is
    // higher than this can touch z = 0 (sqrt((h/2)^2 + r^2), a hair more): the plane response is only evaluated below it
    T k_lin, k_ang, plane_reach;
    T prop_x[4], prop_y[4], pyb_mass, pyb_jx, pyb_jy, pyb_jz, lin_damp, ang_damp, coll_r, coll_h;
    T dt, ctrl_dt, g;
    T pid_gravity, pid_kf;                // DSLPIDControl's own URDF constants (control/BaseControl.py:35-39)
    T hover_lim;                          // 1.5 (HoverAviary.py:108) or 2.0 (MultiHoverAviary.py:121)
    T init_xyz[ADRP_MAX_DRONES][3], init_rpy[ADRP_MAX_DRONES][3], init_quat[ADRP_MAX_DRONES][4], init_vel[ADRP_MAX_DRONES][3],
        init_pqr[ADRP_MAX_DRONES][3], target[ADRP_MAX_DRONES][3];
    // adrp_step_n with ADRP_ACTIONS_PHILOX: every action component is drawn in the kernel, a = u11(Philox4x32-10(key = rng_lo, rng_hi;
    // counter = (global env, 192 + drone, rng_step0 + control step of this launch, 0))): the draw of a random policy
    // (action_space.sample() over Box(-1, 1), envs/BaseRLAviary.py:129-135, envs/MultiRaceAviary.py:284-287)
    int act_rng;
    unsigned rng_lo, rng_hi, rng_step0;
    unsigned long long env_offset;        // global index of env 0 of this shard
};

// Race env constants (envs/MultiRaceAviary.py; config/level*.yaml), by value next to KParams.
constexpr int kGateFields = 6;           // x y z yaw cos(yaw) sin(yaw) of the ACTUAL (randomised) gate
constexpr int kObstFields = 3;           // x y z of the actual obstacle
template <typename T>
struct RaceParams {
    int G, NO;                            // gates, obstacles
    int TF;                               // track fields per env = kGateFields*G + kObstFields*NO
    int R;                                // uniform draws per env reset = 3G + 2NO + 10M
    int compete;                          // RaceMode.COMPETE
    int wrappers;                         // ADRP_WRAP_* (utils/wrapper.py fused into the step)
    int rnd_track, rnd_state, rnd_inertia, disturbances;
    int lut_ticks;                        // length (bits) of the sub-rate schedule LUT
    unsigned seed_lo, seed_hi;            // Philox key
    unsigned long long env_offset;        // global index of env 0 of this shard
    T gate_nom[ADRP_MAX_GATES][4];        // nominal x y z yaw
    int gate_type[ADRP_MAX_GATES];
    T obst_nom[ADRP_MAX_OBSTACLES][3];
    T bounds_hi[3], contact_margin, visibility;
    T gate_lo, gate_hi, obst_lo, obst_hi, pos_lo[3], pos_hi[3], rot_lo[3], rot_hi[3];
    T inertia_lo[4], inertia_hi[4];       // mass, Ixx, Iyy, Izz offsets
    T noise_std, wind_lo[3], wind_hi[3];
    T mel_prev_rpy[ADRP_MAX_DRONES][3];   // init_rpy * pi/180 (MellingerControl.reset, :99-150)
    float mel_mass;                       // firmware g_vehicleMass
    float lpf_b0, lpf_b1, lpf_b2, lpf_a1, lpf_a2;   // gyro lpf2p coefficients (host: filter.c lpf2pSetCutoffFreq)
};

// Mellinger wrapper + firmware state per drone, SoA like the rigid-body state.
enum MelT { MT_PRPY0 = 0, MT_PRPY1, MT_PRPY2, MT_PVEL0, MT_PVEL1, MT_PVEL2, MT_RPM0, MT_RPM1, MT_RPM2, MT_RPM3, MT_COUNT };
enum MelF { MF_GD1X = 0, MF_GD1Y, MF_GD1Z, MF_GD2X, MF_GD2Y, MF_GD2Z, MF_IEX, MF_IEY, MF_IEZ, MF_IMX, MF_IMY, MF_IMZ,
            MF_POR, MF_POP, MF_PSR, MF_PSP, MF_THRUST, MF_ACCZ, MF_COUNT };
enum MelI { MI_TICK = 0, MI_TUMBLE, MI_LAST_ATT, MI_CTL_RP, MI_CTL_Y_ERR, MI_COUNT };   // CTL_RP = roll | pitch<<16; CTL_Y_ERR = yaw | error<<16

constexpr int kHlWordsMax = 80;          // >= sizeof(HlState) / 4 (adrp_planner.cuh): what the host allocates per padded drone
template <typename T>
struct RaceBuffers {
    T* track;               // [TF][N]      actual gates / obstacles of each env
    T* body;                // [4][Dpad]    Bullet mass, Ixx, Iyy, Izz of each drone (_drone_init, :409-431)
    T* mel_t;               // [MT_COUNT][Dpad]
    float* mel_f;           // [MF_COUNT][Dpad]
    int32_t* mel_i;         // [MI_COUNT][Dpad]
    int32_t* gate_latch;    // [Dpad]       current_gate | latch<<16
    int32_t* episode;       // [N]          resets so far (Philox counter word)
    int32_t* shape_gate;    // [N]          RewardWrapper.current_gate_id, -1 = read it on the next step (wrapper.py:96-101)
    T* shape_vec;           // [6][N]       RewardWrapper.current_target xyz, previous_pos xyz
    uint32_t* hl;           // [kHlWords][Dpad] high-level commander state (HlState, word by word); ADRP_ACT_MEL_CMD handles only
    const uint32_t* lut;    // [lut_ticks/32] bit t: (t/500 - (t-1)/500 > 0.002) in float64 (MellingerControl.py:395-411)
};

// Device buffers owned by the handle.
template <typename T>
struct KBuffers {
    T* state;             // [F_COUNT][Dpad]
    T* pid;               // [9][Dpad]   DSLPID integral_pos_e3, integral_rpy_e3, last_rpy3
    T* ring;              // [B*A][Dpad] action buffer of the RL aviaries, ring-indexed by ring_head
    int32_t* step_counter;  // [N] pyb steps since reset (envs/BaseAviary.py:386)
    int32_t* ring_head;     // [N] slot holding the OLDEST action
};

}  // namespace adrp
