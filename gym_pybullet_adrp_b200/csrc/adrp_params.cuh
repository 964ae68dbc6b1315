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
    int pyb_freq;
    int trunc_steps;                      // smallest step_counter with step_counter/PYB_FREQ > EPISODE_LEN_SEC
    // BaseAviary constants (envs/BaseAviary.py:74-128)
    T mass, arm_x, kf, km, jx, jy, jz, jinvx, jinvy, jinvz, gravity, hover_rpm, max_rpm;
    T gnd_eff_coeff, prop_radius, gnd_eff_h_clip, drag_xy, drag_z, dw1, dw2, dw3, speed_limit;
    T prop_x[4], prop_y[4], pyb_mass, pyb_jx, pyb_jy, pyb_jz, lin_damp, ang_damp, coll_r, coll_h;
    T dt, ctrl_dt, g;
    T pid_gravity, pid_kf;                // DSLPIDControl's own URDF constants (control/BaseControl.py:35-39)
    T hover_lim;                          // 1.5 (HoverAviary.py:108) or 2.0 (MultiHoverAviary.py:121)
    T init_xyz[ADRP_MAX_DRONES][3], init_quat[ADRP_MAX_DRONES][4], init_vel[ADRP_MAX_DRONES][3],
        init_pqr[ADRP_MAX_DRONES][3], target[ADRP_MAX_DRONES][3];
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
