/* TEST INFRASTRUCTURE -- C entry points of the CPU oracle (see adrp_oracle_impl.h for the arithmetic).
 *
 * Mirrors the product's C ABI (include/adrp.h) with HOST buffers so that tests can drive the CUDA
 * library and the oracle with the same inputs.  Build: oracle/Makefile  ->  oracle/liboracle.so
 *   gcc -O2 -ffp-contract=off -fno-fast-math -fopenmp -fPIC -shared
 * The product never links or loads this file.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include "../include/adrp.h"

#define REAL double
#define SFX f64
#define SIN sin
#define COS cos
#define SQRT sqrt
#define FABS fabs
#define ATAN2 atan2
#define ASIN asin
#define EXP exp
#define LOG log
#define TAN tan
/* libm float functions of the firmware core in the parity (f64) build: correctly rounded, so that the
 * CUDA fp64 build can reproduce them bit for bit through its own double-precision functions */
#define FW_ATAN2F(y, x) ((float)atan2((double)(y), (double)(x)))
#define FW_SINF(x) ((float)sin((double)(x)))
#define FW_COSF(x) ((float)cos((double)(x)))
#include "adrp_oracle_impl.h"
#include "adrp_oracle_race.h"
#include "adrp_oracle_api.h"
#undef REAL
#undef SFX
#undef SIN
#undef COS
#undef SQRT
#undef FABS
#undef ATAN2
#undef ASIN
#undef EXP
#undef LOG
#undef TAN
#undef FW_ATAN2F
#undef FW_SINF
#undef FW_COSF
#undef CAT2
#undef CAT
#undef FN

#define REAL float
#define SFX f32
#define SIN sinf
#define COS cosf
#define SQRT sqrtf
#define FABS fabsf
#define ATAN2 atan2f
#define ASIN asinf
#define EXP expf
#define LOG logf
#define TAN tanf
#define FW_ATAN2F(y, x) atan2f((y), (x))
#define FW_SINF(x) sinf(x)
#define FW_COSF(x) cosf(x)
#include "adrp_oracle_impl.h"
#include "adrp_oracle_race.h"
#include "adrp_oracle_api.h"

typedef struct orc_any { int precision; void* impl; } orc_any;

int orc_create(const adrp_config* cfg, void** out)
{
    if (!cfg || cfg->abi_version != ADRP_ABI_VERSION) return ADRP_EABI;
    orc_any* a = (orc_any*)calloc(1, sizeof(orc_any));
    a->precision = cfg->precision;
    a->impl = cfg->precision == ADRP_F64 ? (void*)create_f64(cfg) : (void*)create_f32(cfg);
    if (!a->impl) { free(a); return ADRP_EINVAL; }
    *out = a;
    return ADRP_OK;
}

#define DISPATCH(call64, call32) (((orc_any*)h)->precision == ADRP_F64 ? call64 : call32)
#define H64 ((orc_f64*)((orc_any*)h)->impl)
#define H32 ((orc_f32*)((orc_any*)h)->impl)

int orc_destroy(void* h) { DISPATCH(destroy_f64(H64), destroy_f32(H32)); free(h); return ADRP_OK; }
int orc_obs_dim(void* h) { return DISPATCH(H64->O, H32->O); }
int orc_action_dim(void* h) { return DISPATCH(H64->A, H32->A); }
int orc_substeps(void* h) { return DISPATCH(H64->S, H32->S); }
int orc_reset_draws_dim(void* h) { return DISPATCH(reset_draws_dim_f64(H64), reset_draws_dim_f32(H32)); }

int orc_reset(void* h, const uint8_t* env_mask, const void* draws, void* obs)
{
    DISPATCH(reset_f64(H64, env_mask, (const double*)draws, (double*)obs),
             reset_f32(H32, env_mask, (const float*)draws, (float*)obs));
    return ADRP_OK;
}

int orc_set_state(void* h, const void* state)
{
    DISPATCH((void)memcpy(H64->state, state, sizeof(double) * 20 * (size_t)H64->N * H64->M),
             (void)memcpy(H32->state, state, sizeof(float) * 20 * (size_t)H32->N * H32->M));
    return ADRP_OK;
}

int orc_get_state(void* h, void* state)
{
    DISPATCH((void)memcpy(state, H64->state, sizeof(double) * 20 * (size_t)H64->N * H64->M),
             (void)memcpy(state, H32->state, sizeof(float) * 20 * (size_t)H32->N * H32->M));
    return ADRP_OK;
}

/* nthreads <= 1: serial.  Otherwise envs are split over OpenMP threads (the timed CPU baseline). */
int orc_step(void* h, const void* action, const void* noise, void* obs, void* reward,
             uint8_t* terminated, uint8_t* truncated, uint8_t* drone_flags, int nthreads)
{
    DISPATCH(step_f64(H64, (const double*)action, (const double*)noise, (double*)obs, (double*)reward, terminated, truncated, drone_flags, nthreads),
             step_f32(H32, (const float*)action, (const float*)noise, (float*)obs, (float*)reward, terminated, truncated, drone_flags, nthreads));
    return ADRP_OK;
}

/* Race env: only the part of the step after the substep loop (gate progress, observation, flags, counters) on
 * the state currently held -- lets a test evaluate the flag logic on the CUDA kernel's own trajectory. */
int orc_race_post(void* h, void* obs, void* reward, uint8_t* terminated, uint8_t* truncated, uint8_t* drone_flags)
{
    if (DISPATCH(H64->race == NULL, H32->race == NULL)) return ADRP_EINVAL;
    if (((orc_any*)h)->precision == ADRP_F64)
        for (int n = 0; n < H64->N; ++n) race_post_env_f64(H64, n, (double*)obs, (double*)reward, terminated, truncated, drone_flags);
    else
        for (int n = 0; n < H32->N; ++n) race_post_env_f32(H32, n, (float*)obs, (float*)reward, terminated, truncated, drone_flags);
    return ADRP_OK;
}

/* DSLPIDControl.computeControl on n rows (known-answer tests against tests/golden/dslpid_kat.npz). */
int orc_dslpid_control(int precision, int drone_model, const adrp_drone_params* drone, double g, int n, double dt,
                       const void* pos, const void* quat, const void* vel, const void* tpos, const void* trpy,
                       const void* tvel, const void* trates, void* ctrl, void* rpm, void* pos_e, void* yaw_e)
{
    if (precision == ADRP_F64)
        dslpid_rows_f64(drone_model, drone, g, n, dt, pos, quat, vel, tpos, trpy, tvel, trates, ctrl, rpm, pos_e, yaw_e);
    else
        dslpid_rows_f32(drone_model, drone, g, n, dt, pos, quat, vel, tpos, trpy, tvel, trates, ctrl, rpm, pos_e, yaw_e);
    return ADRP_OK;
}

/* ---- firmware entry points for oracle/refstub/firmware_stub.py (lets the reference's own
 * MellingerControl wrapper run against the declared firmware restatement) ---- */
void fw_lpf2p_init(float coef[5], float fs, float fc)
{
    lpf_coef_f32 c;
    lpf2p_init_f32(&c, fs, fc);
    coef[0] = c.b0; coef[1] = c.b1; coef[2] = c.b2; coef[3] = c.a1; coef[4] = c.a2;
}

float fw_lpf2p_apply(const float coef[5], float delay[2], float sample)
{
    lpf_coef_f32 c = {coef[0], coef[1], coef[2], coef[3], coef[4]};
    return lpf2p_apply_f32(&c, &delay[0], &delay[1], sample);
}

/* statics[10] = i_error_x,y,z, i_error_m_x,y,z, prev_omega_roll,pitch, prev_setpoint_omega_roll,pitch
 * sp[16] = pos3 vel3 acc3 rate_deg3 quat4 ; out[4] = roll pitch yaw thrust */
void fw_controller_mellinger(float statics[10], const float sp[16], const float pos[3], const float vel[3],
                             const float q[4], const float gyro_deg[3], float mass, float out[4])
{
    mel_state_f64 m;                 /* the pinned (parity-build) variant; the arithmetic is float32 either way */
    memset(&m, 0, sizeof(m));
    m.i_error_x = statics[0]; m.i_error_y = statics[1]; m.i_error_z = statics[2];
    m.i_error_m_x = statics[3]; m.i_error_m_y = statics[4]; m.i_error_m_z = statics[5];
    m.prev_omega_roll = statics[6]; m.prev_omega_pitch = statics[7];
    m.prev_setpoint_omega_roll = statics[8]; m.prev_setpoint_omega_pitch = statics[9];
    memcpy(m.sp_pos, sp, 12); memcpy(m.sp_vel, sp + 3, 12); memcpy(m.sp_acc, sp + 6, 12);
    memcpy(m.sp_rate_deg, sp + 9, 12); memcpy(m.sp_quat, sp + 12, 16);
    controller_mellinger_f64(&m, mass, pos, vel, q, gyro_deg);
    statics[0] = m.i_error_x; statics[1] = m.i_error_y; statics[2] = m.i_error_z;
    statics[3] = m.i_error_m_x; statics[4] = m.i_error_m_y; statics[5] = m.i_error_m_z;
    statics[6] = m.prev_omega_roll; statics[7] = m.prev_omega_pitch;
    statics[8] = m.prev_setpoint_omega_roll; statics[9] = m.prev_setpoint_omega_pitch;
    out[0] = (float)m.ctl_roll; out[1] = (float)m.ctl_pitch; out[2] = (float)m.ctl_yaw; out[3] = m.ctl_thrust;
}

/* sp[16] as above plus the high-level commander's form of the yaw setpoint (mode.yaw == modeAbs, attitude.yaw in degrees) */
void fw_controller_mellinger_yaw(float statics[10], const float sp[16], int yaw_abs, float yaw_deg, const float pos[3],
                                 const float vel[3], const float q[4], const float gyro_deg[3], float mass, float out[4])
{
    mel_state_f64 m;
    memset(&m, 0, sizeof(m));
    m.i_error_x = statics[0]; m.i_error_y = statics[1]; m.i_error_z = statics[2];
    m.i_error_m_x = statics[3]; m.i_error_m_y = statics[4]; m.i_error_m_z = statics[5];
    m.prev_omega_roll = statics[6]; m.prev_omega_pitch = statics[7];
    m.prev_setpoint_omega_roll = statics[8]; m.prev_setpoint_omega_pitch = statics[9];
    memcpy(m.sp_pos, sp, 12); memcpy(m.sp_vel, sp + 3, 12); memcpy(m.sp_acc, sp + 6, 12);
    memcpy(m.sp_rate_deg, sp + 9, 12); memcpy(m.sp_quat, sp + 12, 16);
    m.sp_yaw_abs = yaw_abs; m.sp_yaw_deg = yaw_deg;
    controller_mellinger_f64(&m, mass, pos, vel, q, gyro_deg);
    statics[0] = m.i_error_x; statics[1] = m.i_error_y; statics[2] = m.i_error_z;
    statics[3] = m.i_error_m_x; statics[4] = m.i_error_m_y; statics[5] = m.i_error_m_z;
    statics[6] = m.prev_omega_roll; statics[7] = m.prev_omega_pitch;
    statics[8] = m.prev_setpoint_omega_roll; statics[9] = m.prev_setpoint_omega_pitch;
    out[0] = (float)m.ctl_roll; out[1] = (float)m.ctl_pitch; out[2] = (float)m.ctl_yaw; out[3] = m.ctl_thrust;
}

/* ---- high-level commander (adrp_oracle_planner.h) for firmware_stub.py: `p` is fw_hl_sizeof() bytes owned by the stub */
int fw_hl_sizeof(void) { return (int)sizeof(planner_f64); }
void fw_hl_init(void* p) { hl_init_f64((planner_f64*)p); }
void fw_hl_stop(void* p) { hl_stop_f64((planner_f64*)p); }
void fw_hl_update_time(void* p, float t) { hl_update_time_f64((planner_f64*)p, t); }
void fw_hl_tell_state(void* p, const float pos[3], const float vel[3], float yaw_deg) { hl_tell_state_f64((planner_f64*)p, pos, vel, yaw_deg); }
int fw_hl_takeoff(void* p, float height, float duration, int use_current_yaw, float yaw) { return hl_takeoff_f64((planner_f64*)p, height, duration, use_current_yaw, yaw); }
int fw_hl_land(void* p, float height, float duration, int use_current_yaw, float yaw) { return hl_land_f64((planner_f64*)p, height, duration, use_current_yaw, yaw); }
int fw_hl_takeoff_vel(void* p, float height, float velocity, int relative) { return hl_takeoff_vel_f64((planner_f64*)p, height, velocity, relative); }
int fw_hl_land_vel(void* p, float height, float velocity, int relative) { return hl_land_vel_f64((planner_f64*)p, height, velocity, relative); }
int fw_hl_go_to(void* p, float x, float y, float z, float yaw, float duration, int relative) { return hl_go_to_f64((planner_f64*)p, x, y, z, yaw, duration, relative); }
int fw_hl_get_setpoint(void* p, const float pos[3], const float vel[3], float yaw_deg, float out[13]) { return hl_get_setpoint_f64((planner_f64*)p, pos, vel, yaw_deg, out); }
