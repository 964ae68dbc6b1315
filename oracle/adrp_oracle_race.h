/* TEST INFRASTRUCTURE -- race env (envs/MultiRaceAviary.py) and Mellinger wrapper
 * (control/MellingerControl.py) restated on the CPU, REAL-generic; plus the DECLARED restatement of
 * the un-vendored firmware core (pycffirmware: controller_mellinger.c, filter.c lpf2p), which is
 * float32 in both builds (its libm calls: FW_ATAN2F / FW_SINF / FW_COSF, see adrp_oracle.c).  PARITY UNPINNED for the firmware core and for the collision / ray queries
 * (oracle/track_geometry.py states the geometric model).  Included by adrp_oracle.c. */

/* ---------------------------------------------------------------- Philox4x32-10 (Salmon et al. 2011) */
static void FN(philox)(uint32_t c0, uint32_t c1, uint32_t c2, uint32_t c3, uint64_t key, uint32_t out[4])
{
    uint32_t k0 = (uint32_t)key, k1 = (uint32_t)(key >> 32);
    for (int r = 0; r < 10; ++r) {
        uint64_t p0 = (uint64_t)0xD2511F53u * c0, p1 = (uint64_t)0xCD9E8D57u * c2;
        uint32_t n0 = (uint32_t)(p1 >> 32) ^ c1 ^ k0, n1 = (uint32_t)p1;
        uint32_t n2 = (uint32_t)(p0 >> 32) ^ c3 ^ k1, n3 = (uint32_t)p0;
        c0 = n0; c1 = n1; c2 = n2; c3 = n3;
        k0 += 0x9E3779B9u; k1 += 0xBB67AE85u;
    }
    out[0] = c0; out[1] = c1; out[2] = c2; out[3] = c3;
}
static REAL FN(u01)(uint32_t x) { return (REAL)(x >> 8) * (REAL)5.9604644775390625e-08; }          /* [0,1) */
static REAL FN(u01_open)(uint32_t x) { return (REAL)((x >> 8) + 1) * (REAL)5.9604644775390625e-08; } /* (0,1] */

#include "adrp_oracle_planner.h"

/* ---------------------------------------------------------------- firmware core (float32, declared) */
typedef struct FN(mel_state) {
    /* wrapper state (MellingerControl.reset, :99-150) */
    REAL prev_rpy[3], prev_vel[3];
    int32_t tick, tumble_counter, last_att_tick, error;
    /* gyro low-pass filters (filter.c lpf2pData): delay_element_1/2 per axis */
    float gyro_d1[3], gyro_d2[3];
    /* controller_mellinger.c file-scope statics */
    float i_error_x, i_error_y, i_error_z, i_error_m_x, i_error_m_y, i_error_m_z;
    float prev_omega_roll, prev_omega_pitch, prev_setpoint_omega_roll, prev_setpoint_omega_pitch;
    /* control_t */
    int16_t ctl_roll, ctl_pitch, ctl_yaw;
    float ctl_thrust;
    /* setpoint_t (full-state command, :510-543) */
    float sp_pos[3], sp_vel[3], sp_acc[3], sp_rate_deg[3], sp_quat[4];
    /* ... as the high-level commander leaves it (crtpCommanderHighLevelGetSetpoint): mode.yaw = modeAbs with
     * attitude.yaw in degrees, mode.quat = modeDisable */
    float sp_yaw_deg;
    int32_t sp_yaw_abs;
    /* high-level commander (MellingerControl.py:292-303,:369-374,:547-699) */
    int32_t hl_active;             /* not full_state_cmd_override */
    FN(planner) plan;
    /* state_t as last written by _update_state (:446-471; C floats): what _notifySetpointStop hands to TellState */
    float st_pos[3], st_vel[3], st_yaw_deg;
    /* motor commands: what MultiRaceAviary keeps in self.rpms (prev_rpms lives in the state row 16:20) */
    REAL rpms[4];
    /* state_t.acc.z as stored (float) for the tumble check (:382) */
    float state_acc_z;
} FN(mel_state);

typedef struct FN(lpf_coef) { float b0, b1, b2, a1, a2; } FN(lpf_coef);

/* filter.c lpf2pSetCutoffFreq */
static void FN(lpf2p_init)(FN(lpf_coef)* c, float sample_freq, float cutoff_freq)
{
    float fr = sample_freq / cutoff_freq;
    float ohm = tanf(3.14159265358979323846f / fr);
    float cc = 1.0f + 2.0f * cosf(3.14159265358979323846f / 4.0f) * ohm + ohm * ohm;
    c->b0 = ohm * ohm / cc;
    c->b1 = 2.0f * c->b0;
    c->b2 = c->b0;
    c->a1 = 2.0f * (ohm * ohm - 1.0f) / cc;
    c->a2 = (1.0f - 2.0f * cosf(3.14159265358979323846f / 4.0f) * ohm + ohm * ohm) / cc;
}

/* filter.c lpf2pApply */
static float FN(lpf2p_apply)(const FN(lpf_coef)* c, float* d1, float* d2, float sample)
{
    float d0 = sample - *d1 * c->a1 - *d2 * c->a2;
    if (!isfinite(d0)) d0 = sample;
    float out = d0 * c->b0 + *d1 * c->b1 + *d2 * c->b2;
    *d2 = *d1;
    *d1 = d0;
    return out;
}

static float FN(clampf)(float v, float lo, float hi) { return v < lo ? lo : (v > hi ? hi : v); }
static float FN(radiansf)(float d) { return (3.14159265358979323846f / 180.0f) * d; }
static float FN(degreesf)(float r) { return (180.0f / 3.14159265358979323846f) * r; }

/* controller_mellinger.c controllerMellinger(), setpoint modes x/y/z/quat = modeAbs, roll/pitch/yaw =
 * modeDisable (what _sendFullStateCmd sets).  Inputs: state position/velocity/attitudeQuaternion,
 * gyro in deg/s.  eR is the matrix form the firmware's expanded "fast version" was generated from. */
static void FN(controller_mellinger)(FN(mel_state)* m, float mass, const float pos[3], const float vel[3],
                                     const float q[4], const float gyro_deg[3])
{
    const float massThrust = 132000.0f;
    const float kp_xy = 0.4f, kd_xy = 0.2f, ki_xy = 0.05f, i_range_xy = 2.0f;
    const float kp_z = 1.25f, kd_z = 0.4f, ki_z = 0.05f, i_range_z = 0.4f;
    const float kR_xy = 70000.0f, kw_xy = 20000.0f, ki_m_xy = 0.0f, i_range_m_xy = 1.0f;
    const float kR_z = 60000.0f, kw_z = 12000.0f, ki_m_z = 500.0f, i_range_m_z = 1500.0f;
    const float kd_omega_rp = 200.0f, GRAVITY_MAGNITUDE = 9.81f;
    const float dt = (float)(1.0f / 500.0f);
    float r_error[3], v_error[3], tt[3];
    for (int k = 0; k < 3; ++k) { r_error[k] = m->sp_pos[k] - pos[k]; v_error[k] = m->sp_vel[k] - vel[k]; }
    m->i_error_z += r_error[2] * dt; m->i_error_z = FN(clampf)(m->i_error_z, -i_range_z, i_range_z);
    m->i_error_x += r_error[0] * dt; m->i_error_x = FN(clampf)(m->i_error_x, -i_range_xy, i_range_xy);
    m->i_error_y += r_error[1] * dt; m->i_error_y = FN(clampf)(m->i_error_y, -i_range_xy, i_range_xy);
    tt[0] = mass * m->sp_acc[0] + kp_xy * r_error[0] + kd_xy * v_error[0] + ki_xy * m->i_error_x;
    tt[1] = mass * m->sp_acc[1] + kp_xy * r_error[1] + kd_xy * v_error[1] + ki_xy * m->i_error_y;
    tt[2] = mass * (m->sp_acc[2] + GRAVITY_MAGNITUDE) + kp_z * r_error[2] + kd_z * v_error[2] + ki_z * m->i_error_z;
    /* desired yaw from the setpoint quaternion (mode.quat == modeAbs): quat2rpy().z in degrees */
    const float sx = m->sp_quat[0], sy = m->sp_quat[1], sz = m->sp_quat[2], sw = m->sp_quat[3];
    float desiredYaw = FN(degreesf)(FW_ATAN2F(2.0f * (sw * sz + sx * sy), 1 - 2 * (sy * sy + sz * sz)));
    if (m->sp_yaw_abs) desiredYaw = m->sp_yaw_deg;                 /* mode.yaw == modeAbs: setpoint->attitude.yaw */
    /* R = quat2rotmat(state quaternion) */
    const float x = q[0], y = q[1], z = q[2], w = q[3];
    float R[3][3] = {{1 - 2 * y * y - 2 * z * z, 2 * x * y - 2 * z * w, 2 * x * z + 2 * y * w},
                     {2 * x * y + 2 * z * w, 1 - 2 * x * x - 2 * z * z, 2 * y * z - 2 * x * w},
                     {2 * x * z - 2 * y * w, 2 * y * z + 2 * x * w, 1 - 2 * x * x - 2 * y * y}};
    float current_thrust = tt[0] * R[0][2] + tt[1] * R[1][2] + tt[2] * R[2][2];
    float nt = sqrtf(tt[0] * tt[0] + tt[1] * tt[1] + tt[2] * tt[2]);
    float zd[3] = {tt[0] / nt, tt[1] / nt, tt[2] / nt};
    float xc[3] = {FW_COSF(FN(radiansf)(desiredYaw)), FW_SINF(FN(radiansf)(desiredYaw)), 0};
    float yd[3] = {zd[1] * xc[2] - zd[2] * xc[1], zd[2] * xc[0] - zd[0] * xc[2], zd[0] * xc[1] - zd[1] * xc[0]};
    float ny = sqrtf(yd[0] * yd[0] + yd[1] * yd[1] + yd[2] * yd[2]);
    yd[0] /= ny; yd[1] /= ny; yd[2] /= ny;
    float xd[3] = {yd[1] * zd[2] - yd[2] * zd[1], yd[2] * zd[0] - yd[0] * zd[2], yd[0] * zd[1] - yd[1] * zd[0]};
    /* eRM = Rdes^T R - R^T Rdes, Rdes = [xd yd zd] (columns); eR = (eRM[2][1], -eRM[0][2], eRM[1][0]) */
    float Rd[3][3] = {{xd[0], yd[0], zd[0]}, {xd[1], yd[1], zd[1]}, {xd[2], yd[2], zd[2]}};
    float e21 = 0, e02 = 0, e10 = 0;
    for (int k = 0; k < 3; ++k) {
        e21 += Rd[k][2] * R[k][1] - R[k][2] * Rd[k][1];
        e02 += Rd[k][0] * R[k][2] - R[k][0] * Rd[k][2];
        e10 += Rd[k][1] * R[k][0] - R[k][1] * Rd[k][0];
    }
    /* the firmware's expanded expression yields vee(eRM) = (eRM[2][1], eRM[0][2], eRM[1][0]) and then flips
     * eR.y ("Account for Crazyflie coordinate system"), which is the commented slow version's -eRM[0][2] */
    float eR[3] = {e21, -e02, e10};
    float rate_roll = FN(radiansf)(gyro_deg[0]), rate_pitch = -FN(radiansf)(gyro_deg[1]), rate_yaw = FN(radiansf)(gyro_deg[2]);
    float ew[3] = {FN(radiansf)(m->sp_rate_deg[0]) - rate_roll, -FN(radiansf)(m->sp_rate_deg[1]) - rate_pitch,
                   FN(radiansf)(m->sp_rate_deg[2]) - rate_yaw};
    float err_d_roll = ((FN(radiansf)(m->sp_rate_deg[0]) - m->prev_setpoint_omega_roll) - (rate_roll - m->prev_omega_roll)) / dt;
    float err_d_pitch = (-(FN(radiansf)(m->sp_rate_deg[1]) - m->prev_setpoint_omega_pitch) - (rate_pitch - m->prev_omega_pitch)) / dt;
    m->prev_omega_roll = rate_roll; m->prev_omega_pitch = rate_pitch;
    m->prev_setpoint_omega_roll = FN(radiansf)(m->sp_rate_deg[0]);
    m->prev_setpoint_omega_pitch = FN(radiansf)(m->sp_rate_deg[1]);
    m->i_error_m_x += (-eR[0]) * dt; m->i_error_m_x = FN(clampf)(m->i_error_m_x, -i_range_m_xy, i_range_m_xy);
    m->i_error_m_y += (-eR[1]) * dt; m->i_error_m_y = FN(clampf)(m->i_error_m_y, -i_range_m_xy, i_range_m_xy);
    m->i_error_m_z += (-eR[2]) * dt; m->i_error_m_z = FN(clampf)(m->i_error_m_z, -i_range_m_z, i_range_m_z);
    float Mx = -kR_xy * eR[0] + kw_xy * ew[0] + ki_m_xy * m->i_error_m_x + kd_omega_rp * err_d_roll;
    float My = -kR_xy * eR[1] + kw_xy * ew[1] + ki_m_xy * m->i_error_m_y + kd_omega_rp * err_d_pitch;
    float Mz = -kR_z * eR[2] + kw_z * ew[2] + ki_m_z * m->i_error_m_z;
    m->ctl_thrust = massThrust * current_thrust;
    if (m->ctl_thrust > 0) {
        m->ctl_roll = (int16_t)FN(clampf)(Mx, -32000, 32000);
        m->ctl_pitch = (int16_t)FN(clampf)(My, -32000, 32000);
        m->ctl_yaw = (int16_t)FN(clampf)(-Mz, -32000, 32000);
    } else {
        m->ctl_roll = m->ctl_pitch = m->ctl_yaw = 0;
        m->i_error_x = m->i_error_y = m->i_error_z = 0;          /* controllerMellingerReset() */
        m->i_error_m_x = m->i_error_m_y = m->i_error_m_z = 0;
    }
}

/* ---------------------------------------------------------------- race env state */
typedef struct FN(orc_race) {
    int G, O;                      /* gates, obstacles */
    REAL* gates_actual;            /* (N, G, 4) x y z yaw  (MultiRaceAviary._addObstacles, :347-403) */
    REAL* obst_actual;             /* (N, O, 3) */
    REAL* body;                    /* (N, M, 4) Bullet mass, Ixx, Iyy, Izz per drone (_drone_init, :409-431) */
    int32_t* current_gate;         /* (N, M)  (:83) */
    uint8_t* latch;                /* (N, M)  ADRP_FLAG_ELIMINATED | ADRP_FLAG_FINISHED */
    int32_t* episode;              /* (N) resets so far: Philox counter word */
    FN(mel_state)* mel;            /* (N, M) */
    REAL* shaping;                 /* (N, 7) RewardWrapper state: current_gate_id (-1: not yet initialised),
                                      current_target xyz, previous_pos xyz (utils/wrapper.py:78-80,98-101) */
    FN(lpf_coef) gyro_lpf;         /* lpf2pInit(gyrolpf, 500, ACCEL_LPF_CUTOFF_FREQ = 30)  (:116-117, E12) */
} FN(orc_race);

static FN(orc_race)* FN(race_create)(FN(orc)* h)
{
    FN(orc_race)* r = (FN(orc_race)*)calloc(1, sizeof(FN(orc_race)));
    r->G = h->cfg.track.num_gates; r->O = h->cfg.track.num_obstacles;
    size_t N = (size_t)h->N, M = (size_t)h->M;
    r->gates_actual = (REAL*)calloc(N * (r->G ? r->G : 1) * 4, sizeof(REAL));
    r->obst_actual = (REAL*)calloc(N * (r->O ? r->O : 1) * 3, sizeof(REAL));
    r->body = (REAL*)calloc(N * M * 4, sizeof(REAL));
    r->current_gate = (int32_t*)calloc(N * M, sizeof(int32_t));
    r->latch = (uint8_t*)calloc(N * M, 1);
    r->episode = (int32_t*)calloc(N, sizeof(int32_t));
    r->mel = (FN(mel_state)*)calloc(N * M, sizeof(FN(mel_state)));
    r->shaping = (REAL*)calloc(N * 7, sizeof(REAL));
    for (size_t n = 0; n < N; ++n) r->shaping[7 * n] = (REAL)-1;
    FN(lpf2p_init)(&r->gyro_lpf, 500.0f, 30.0f);
    h->O = 12 + 5 * r->G + 4 * r->O + 1 + (h->cfg.race_mode == ADRP_RACE_COMPETE ? 6 * (h->M - 1) : 0);
    return r;
}

static void FN(race_destroy)(FN(orc_race)* r)
{
    free(r->gates_actual); free(r->obst_actual); free(r->body); free(r->current_gate); free(r->latch);
    free(r->episode); free(r->mel); free(r->shaping); free(r);
}

/* uniform draws consumed by one reset: gates 3 each, obstacles 2 each, per drone 4 inertia + 3 pos + 3 rot */
static int FN(race_draws_dim)(const FN(orc)* h)
{
    return 3 * h->cfg.track.num_gates + 2 * h->cfg.track.num_obstacles + 10 * h->M;
}

static REAL FN(map_range)(REAL u, const double r[2]) { return (REAL)r[0] + u * ((REAL)r[1] - (REAL)r[0]); }

/* MultiRaceAviary.reset (:127-167): track randomisation, _drone_init, controller reset.
 * draws == NULL: Philox stream (seed; global env; episode).  Draw order = race_draws_dim layout. */
static void FN(race_init_env)(FN(orc)* h, int n, const REAL* draws)
{
    FN(orc_race)* r = h->race;
    const adrp_config* c = &h->cfg;
    const int M = h->M, R = FN(race_draws_dim)(h);
    REAL u[3 * ADRP_MAX_GATES + 2 * ADRP_MAX_OBSTACLES + 10 * ADRP_MAX_DRONES];
    r->shaping[7 * (size_t)n] = (REAL)-1;                             /* RewardWrapper.reset: re-read on the next step */
    if (draws) {
        for (int k = 0; k < R; ++k) u[k] = draws[k];
    } else {
        for (int k = 0; k < R; k += 4) {
            uint32_t o[4];
            FN(philox)((uint32_t)(c->env_offset + (uint64_t)n), 0u, (uint32_t)(k / 4), (uint32_t)r->episode[n], c->seed, o);
            for (int j = 0; j < 4 && k + j < R; ++j) u[k + j] = FN(u01)(o[j]);
        }
    }
    r->episode[n] += 1;
    int d = 0;
    for (int g = 0; g < r->G; ++g) {                                /* :359-368 */
        REAL* ga = r->gates_actual + ((size_t)n * r->G + g) * 4;
        for (int k = 0; k < 4; ++k) ga[k] = (REAL)c->track.gates[g][k];
        if (c->rnd.random_gates_obstacles) {
            ga[0] += FN(map_range)(u[d], c->rnd.gate_range);
            ga[1] += FN(map_range)(u[d + 1], c->rnd.gate_range);
            ga[3] += FN(map_range)(u[d + 2], c->rnd.gate_range);
        }
        d += 3;
    }
    for (int o = 0; o < r->O; ++o) {                                /* :370-380 */
        REAL* oa = r->obst_actual + ((size_t)n * r->O + o) * 3;
        for (int k = 0; k < 3; ++k) oa[k] = (REAL)c->track.obstacles[o][k];
        if (c->rnd.random_gates_obstacles) {
            oa[0] += FN(map_range)(u[d], c->rnd.obstacle_range);
            oa[1] += FN(map_range)(u[d + 1], c->rnd.obstacle_range);
        }
        d += 2;
    }
    /* _drone_init (:407-467): the property dict is mutated in the loop, so offsets ACCUMULATE over drones (E16) */
    REAL prop[4] = {(REAL)c->drone.pyb_mass, (REAL)c->drone.pyb_ixx, (REAL)c->drone.pyb_iyy, (REAL)c->drone.pyb_izz};
    const double* ranges[4] = {c->rnd.mass_range, c->rnd.ixx_range, c->rnd.iyy_range, c->rnd.izz_range};
    for (int i = 0; i < M; ++i) {
        REAL* st = h->state + ((size_t)n * M + i) * 20;
        if (c->rnd.random_drone_inertia)
            for (int k = 0; k < 4; ++k) prop[k] = FN(clip)(prop[k] + FN(map_range)(u[d + k], ranges[k]), 0, 100);   /* :419-424 */
        d += 4;
        for (int k = 0; k < 4; ++k) r->body[((size_t)n * M + i) * 4 + k] = prop[k];
        REAL e[3] = {(REAL)c->init_rpy[i][0], (REAL)c->init_rpy[i][1], (REAL)c->init_rpy[i][2]};
        if (c->rnd.random_drone_state) {
            for (int k = 0; k < 3; ++k) st[k] += FN(map_range)(u[d + k], c->rnd.pos_range[k]);       /* :434-450 */
            for (int k = 0; k < 3; ++k) e[k] += FN(map_range)(u[d + 3 + k], c->rnd.rot_range[k]);
        }
        d += 6;
        FN(quat_from_euler)(e, st + 3);                              /* :456-461 (rpy used without deg->rad) */
        FN(pose_round_trip)(st + 3);
        /* MellingerControl.reset(initial_obs) (:99-150): initial_obs predates _drone_init (E17) */
        FN(mel_state)* m = r->mel + (size_t)n * M + i;
        memset(m, 0, sizeof(*m));
        for (int k = 0; k < 3; ++k) { m->prev_rpy[k] = (REAL)(c->init_rpy[i][k] * (M_PI / 180.0)); m->prev_vel[k] = 0; }
        m->state_acc_z = 1.0f;                                       /* _update_state(0, pos, vel, VEC3_UP, rpy) */
        {   /* crtpCommanderHighLevelInit(); ...TellState(self.state): the NOMINAL initial pose (initial_obs, E17) */
            FN(hl_init)(&m->plan);
            for (int k = 0; k < 3; ++k) { m->st_pos[k] = (float)(REAL)c->init_xyz[i][k]; m->st_vel[k] = 0; }
            m->st_yaw_deg = (float)(m->prev_rpy[2] * (REAL)(180.0 / M_PI));
            FN(hl_tell_state)(&m->plan, m->st_pos, m->st_vel, m->st_yaw_deg);
        }
        r->current_gate[(size_t)n * M + i] = 0;
        r->latch[(size_t)n * M + i] = 0;
    }
}

/* ---------------------------------------------------------------- MellingerControl.computeControl (:154-262) */
static void FN(mellinger_tick)(FN(orc)* h, FN(mel_state)* m, const REAL* st, const REAL noise4[4])
{
    const REAL FIRMWARE_DT = (REAL)(1.0 / 500), RAD_TO_DEG = (REAL)(180.0 / M_PI), DEG_TO_RAD = (REAL)(M_PI / 180.0);
    REAL rpy[3];
    FN(euler)(st + 3, rpy);                                          /* obs[i, 3:6] (:242) */
    REAL rates[3], acc[3];
    for (int k = 0; k < 3; ++k) {
        rates[k] = (rpy[k] - m->prev_rpy[k]) / FIRMWARE_DT;          /* :209 */
        m->prev_rpy[k] = rpy[k];
        acc[k] = (st[7 + k] - m->prev_vel[k]) / FIRMWARE_DT / (REAL)9.8 + (k == 2 ? (REAL)1 : (REAL)0);   /* :213 */
        m->prev_vel[k] = st[7 + k];
    }
    /* _update_state (:446-471): struct fields are C floats */
    REAL rd[3] = {rpy[0] * RAD_TO_DEG * DEG_TO_RAD, rpy[1] * RAD_TO_DEG * DEG_TO_RAD, rpy[2] * RAD_TO_DEG * DEG_TO_RAD};
    /* utils.get_quaternion_from_euler (utils/utils.py:20-43) */
    REAL cr = COS(rd[0] / 2), sr = SIN(rd[0] / 2), cp = COS(rd[1] / 2), sp = SIN(rd[1] / 2), cy = COS(rd[2] / 2), sy = SIN(rd[2] / 2);
    float q[4] = {(float)(sr * cp * cy - cr * sp * sy), (float)(cr * sp * cy + sr * cp * sy),
                  (float)(cr * cp * sy - sr * sp * cy), (float)(cr * cp * cy + sr * sp * sy)};
    float pos[3] = {(float)st[0], (float)st[1], (float)st[2]}, vel[3] = {(float)st[7], (float)st[8], (float)st[9]};
    m->state_acc_z = (float)acc[2];
    /* _update_sensorData (:347-365): gyro deg/s through lpf2p (the acc filters feed nothing the controller reads) */
    float gyro[3];
    for (int k = 0; k < 3; ++k)
        gyro[k] = FN(lpf2p_apply)(&h->race->gyro_lpf, &m->gyro_d1[k], &m->gyro_d2[k], (float)(rates[k] * RAD_TO_DEG));
    /* _update_setpoint (:369-374): while no full-state command overrides it, the high-level commander writes the setpoint.
     * state_t fields are C floats: position, velocity, attitude.yaw = yaw in degrees (_update_state, :446-471) */
    for (int k = 0; k < 3; ++k) { m->st_pos[k] = pos[k]; m->st_vel[k] = vel[k]; }
    m->st_yaw_deg = (float)(rpy[2] * RAD_TO_DEG);
    if (m->hl_active) {
        float sp13[13];
        const float yaw_deg = m->st_yaw_deg;
        FN(hl_tell_state)(&m->plan, pos, vel, yaw_deg);
        FN(hl_update_time)(&m->plan, (float)((double)m->tick / 500.0));
        if (FN(hl_get_setpoint)(&m->plan, pos, vel, yaw_deg, sp13)) {
            for (int k = 0; k < 3; ++k) { m->sp_pos[k] = sp13[k]; m->sp_vel[k] = sp13[3 + k]; m->sp_acc[k] = sp13[6 + k]; m->sp_rate_deg[k] = sp13[9 + k]; }
            m->sp_yaw_deg = sp13[12];
            m->sp_yaw_abs = 1;
        }
    }
    /* _step_controller (:378-419) */
    REAL pwms[4] = {0, 0, 0, 0};
    if (m->state_acc_z < -0.5f) m->tumble_counter += 1; else m->tumble_counter = 0;   /* :382-385 */
    if (m->tumble_counter >= 30) {
        m->tick += 1; m->error = 1;                                  /* :387-391: zero PWM */
    } else {
        /* (cur_time - last_att_pid_call > 0.002) in float64, as Python evaluates it (:395-411); _tick 0 and 2
         * both execute the whole Mellinger law (RATE_DO_EXECUTE(ATTITUDE_RATE, tick)), _tick 1 holds `control` */
        double cur_time = (double)m->tick / 500.0, last = (double)m->last_att_tick / 500.0;
        if (cur_time - last > 0.002) {
            m->last_att_tick = m->tick;
            FN(controller_mellinger)(m, (float)h->cfg.mel_mass, pos, vel, q, gyro);
        }
        m->tick += 1;
        /* _compute_pwms (:423-442), float64 */
        REAL r = (REAL)m->ctl_roll / 2, p = (REAL)m->ctl_pitch / 2, y = (REAL)m->ctl_yaw, t = (REAL)m->ctl_thrust;
        REAL th[4] = {t - r + p + y, t - r - p - y, t + r - p + y, t + r + p - y};
        for (int k = 0; k < 4; ++k) {
            REAL x = FN(clip)(th[k], 0, 65535) / (REAL)65535 * (REAL)60;
            REAL volts = (REAL)-0.0006239 * (x * x) + (REAL)0.088 * x;
            REAL pct = volts / (REAL)3 < (REAL)1 ? volts / (REAL)3 : (REAL)1;
            pwms[k] = pct * (REAL)65535;
        }
    }
    /* :246-262 */
    const REAL KF = (REAL)3.16e-10, SCALE = (REAL)0.2685, CONST = (REAL)4070.3;
    REAL thrust[4];
    for (int k = 0; k < 4; ++k) {
        REAL c = FN(clip)(pwms[k], 20000, 65535);
        REAL rp = SCALE * c + CONST;
        thrust[k] = KF * (rp * rp);
    }
    for (int k = 0; k < 4; ++k) {
        REAL tk = thrust[3 - k] + noise4[k];                         /* thrust[[3,2,1,0]] + disturbance (:250-253) */
        if (tk < 0) tk = 0;                                          /* _thr2pwm (:307-343) */
        REAL pwm = (SQRT(tk / (REAL)1 / KF) - CONST) / SCALE;
        pwm = FN(clip)(pwm, 20000, 65535);
        m->rpms[k] = SCALE * pwm + CONST;
    }
}

/* ---------------------------------------------------------------- geometry (oracle/track_geometry.py in C) */
static REAL FN(gap_circle_rect)(REAL px, REAL py, REAL r, REAL cx, REAL cy, REAL yaw, REAL hx, REAL hy)
{
    REAL c = COS(yaw), s = SIN(yaw), dx = px - cx, dy = py - cy;
    REAL lx = c * dx + s * dy, ly = -s * dx + c * dy;
    REAL ex = FABS(lx) - hx, ey = FABS(ly) - hy;
    if (ex > 0 || ey > 0) {
        REAL ox = ex > 0 ? ex : 0, oy = ey > 0 ? ey : 0;
        return SQRT(ox * ox + oy * oy) - r;
    }
    return (ex > ey ? ex : ey) - r;
}
static REAL FN(gap_circle_circle)(REAL px, REAL py, REAL r, REAL cx, REAL cy, REAL r2)
{
    REAL dx = px - cx, dy = py - cy;
    return SQRT(dx * dx + dy * dy) - r - r2;
}
static REAL FN(combine)(REAL dxy, REAL dz)
{
    REAL a = dxy > 0 ? dxy : 0, b = dz > 0 ? dz : 0;
    if (a == 0 && b == 0) return dxy > dz ? dxy : dz;
    return SQRT(a * a + b * b);
}
static REAL FN(prism_rect)(REAL px, REAL py, REAL pz, REAL hz, REAL r, REAL cx, REAL cy, REAL yaw, REAL hx, REAL hy, REAL cz, REAL hz2)
{
    return FN(combine)(FN(gap_circle_rect)(px, py, r, cx, cy, yaw, hx, hy), FABS(pz - cz) - hz - hz2);
}
static REAL FN(prism_circ)(REAL px, REAL py, REAL pz, REAL hz, REAL r, REAL cx, REAL cy, REAL r2, REAL cz, REAL hz2)
{
    return FN(combine)(FN(gap_circle_circle)(px, py, r, cx, cy, r2), FABS(pz - cz) - hz - hz2);
}
static REAL FN(drone_hz)(const FN(orc)* h, const REAL* st)
{
    REAL R[3][3];
    FN(rotmat)(st + 3, R);
    REAL s2 = (REAL)1 - R[2][2] * R[2][2];
    return (REAL)0.5 * h->D.coll_h * FABS(R[2][2]) + h->D.coll_r * SQRT(s2 > 0 ? s2 : 0);
}
/* distance drone <-> gate g (portal.urdf / low_portal.urdf shapes as vertical prisms) */
static REAL FN(gate_distance)(const FN(orc)* h, const REAL* st, REAL hz, const REAL* ga, int low)
{
    const REAL r = h->D.coll_r, x = ga[0], y = ga[1], z = ga[2], yaw = ga[3];
    const REAL c = COS(yaw), s = SIN(yaw);
    REAL d = FN(prism_rect)(st[0], st[1], st[2], hz, r, x, y, yaw, (REAL)0.25, (REAL)0.025, z - (REAL)0.225, (REAL)0.025), e;
    e = FN(prism_rect)(st[0], st[1], st[2], hz, r, x, y, yaw, (REAL)0.25, (REAL)0.025, z + (REAL)0.225, (REAL)0.025); if (e < d) d = e;
    e = FN(prism_rect)(st[0], st[1], st[2], hz, r, x + (REAL)0.225 * c, y + (REAL)0.225 * s, yaw, (REAL)0.025, (REAL)0.025, z, (REAL)0.25); if (e < d) d = e;
    e = FN(prism_rect)(st[0], st[1], st[2], hz, r, x - (REAL)0.225 * c, y - (REAL)0.225 * s, yaw, (REAL)0.025, (REAL)0.025, z, (REAL)0.25); if (e < d) d = e;
    if (low) e = FN(prism_rect)(st[0], st[1], st[2], hz, r, x, y, yaw, (REAL)0.075, (REAL)0.075, z - (REAL)0.4, (REAL)0.125);
    else e = FN(prism_circ)(st[0], st[1], st[2], hz, r, x, y, (REAL)0.05, z - (REAL)0.6, (REAL)0.4);
    return e < d ? e : d;
}
static REAL FN(obstacle_distance)(const FN(orc)* h, const REAL* st, REAL hz, const REAL* oa)
{
    const REAL r = h->D.coll_r;
    REAL d = FN(prism_circ)(st[0], st[1], st[2], hz, r, oa[0], oa[1], (REAL)0.05, oa[2], (REAL)0.4);
    REAL e = FN(prism_rect)(st[0], st[1], st[2], hz, r, oa[0], oa[1], 0, (REAL)0.075, (REAL)0.075, oa[2] - (REAL)0.4, (REAL)0.125);
    return e < d ? e : d;
}

/* First hit fraction of segment o + t d (t in [0,1]) with the solid oriented cylinder; -1 = no hit
 * (also when the segment starts inside: Bullet's convex ray cast reports nothing from inside). */
static REAL FN(ray_cylinder)(const REAL o[3], const REAL d[3], const REAL c[3], const REAL ax[3], REAL r, REAL hh)
{
    REAL oc[3] = {o[0] - c[0], o[1] - c[1], o[2] - c[2]};
    REAL oa = oc[0] * ax[0] + oc[1] * ax[1] + oc[2] * ax[2];
    REAL da = d[0] * ax[0] + d[1] * ax[1] + d[2] * ax[2];
    REAL op[3], dp[3];
    for (int k = 0; k < 3; ++k) { op[k] = oc[k] - oa * ax[k]; dp[k] = d[k] - da * ax[k]; }
    REAL rad2 = op[0] * op[0] + op[1] * op[1] + op[2] * op[2];
    if (FABS(oa) <= hh && rad2 <= r * r) return -1;
    REAL t0 = 0, t1 = 1;
    if (da == 0) {
        if (FABS(oa) > hh) return -1;
    } else {
        REAL ta = (-hh - oa) / da, tb = (hh - oa) / da;
        if (ta > tb) { REAL tmp = ta; ta = tb; tb = tmp; }
        if (ta > t0) t0 = ta;
        if (tb < t1) t1 = tb;
        if (t0 > t1) return -1;
    }
    REAL A = dp[0] * dp[0] + dp[1] * dp[1] + dp[2] * dp[2];
    REAL B = op[0] * dp[0] + op[1] * dp[1] + op[2] * dp[2];
    REAL C = rad2 - r * r;
    if (A == 0) {
        if (C > 0) return -1;
    } else {
        REAL disc = B * B - A * C;
        if (disc < 0) return -1;
        REAL sq = SQRT(disc), ta = (-B - sq) / A, tb = (-B + sq) / A;
        if (ta > t0) t0 = ta;
        if (tb < t1) t1 = tb;
        if (t0 > t1) return -1;
    }
    return t0;
}

/* MultiRaceAviary._gate_progress (:471-506) for every drone of env n */
static void FN(gate_progress)(FN(orc)* h, int n)
{
    FN(orc_race)* r = h->race;
    const int M = h->M;
    for (int i = 0; i < M; ++i) {
        int32_t* cg = r->current_gate + (size_t)n * M + i;
        const int gate = *cg;
        if (r->G > 0 && gate < r->G) {
            const REAL* ga = r->gates_actual + ((size_t)n * r->G + gate) * 4;
            const REAL height = h->cfg.track.gate_type[gate] == 0 ? (REAL)1.0 : (REAL)0.525;   /* :479-482 */
            const REAL half_length = (REAL)0.1875;
            const REAL delta_x = (REAL)0.05 * COS(ga[3]), delta_y = (REAL)0.05 * SIN(ga[3]);
            int passed = 0;
            for (int k = 0; k < 7 && !passed; ++k) {
                /* ray order of the reference: centre, +1, -1, +2, -2, +3, -3 */
                const int mult = k == 0 ? 0 : ((k + 1) / 2) * ((k & 1) ? 1 : -1);
                REAL fr[3] = {ga[0] + mult * delta_x, ga[1] + mult * delta_y, height - half_length};
                REAL dir[3] = {0, 0, (height + half_length) - (height - half_length)};
                REAL best = -1; int best_id = -1;
                for (int j = 0; j < M; ++j) {                      /* first body hit along the ray */
                    const REAL* sj = h->state + ((size_t)n * M + j) * 20;
                    REAL R[3][3];
                    FN(rotmat)(sj + 3, R);
                    REAL ax[3] = {R[0][2], R[1][2], R[2][2]};
                    REAL t = FN(ray_cylinder)(fr, dir, sj, ax, h->D.coll_r, (REAL)0.5 * h->D.coll_h);
                    if (t >= 0 && (best < 0 || t < best)) { best = t; best_id = j; }
                }
                if (best_id == i && best < (REAL)0.9999) passed = 1;   /* :502 */
            }
            if (passed) *cg += 1;
        }
        if (gate >= r->G) r->latch[(size_t)n * M + i] |= ADRP_FLAG_FINISHED;   /* :505-506 (one step late, E18) */
    }
}

/* MultiRaceAviary._computeObs (:566-661) */
static void FN(race_write_obs)(FN(orc)* h, int n, REAL* obs)
{
    FN(orc_race)* r = h->race;
    const adrp_config* c = &h->cfg;
    const int M = h->M, G = r->G, NO = r->O;
    const REAL vis = (REAL)c->track.visibility_range;
    for (int i = 0; i < M; ++i) {
        const REAL* st = h->state + ((size_t)n * M + i) * 20;
        REAL* o = obs + ((size_t)n * M + i) * h->O;
        const REAL hz = FN(drone_hz)(h, st);
        FN(write_kin12)(st, o);
        int k = 12;
        for (int g = 0; g < G; ++g) {
            const REAL* ga = r->gates_actual + ((size_t)n * G + g) * 4;
            const int in_range = FN(gate_distance)(h, st, hz, ga, c->track.gate_type[g]) < vis;    /* :614-626 */
            for (int q = 0; q < 4; ++q) o[k + 4 * g + q] = in_range ? ga[q] : (REAL)c->track.gates[g][q];
            o[k + 4 * G + g] = (REAL)in_range;
        }
        k += 5 * G;
        for (int b = 0; b < NO; ++b) {
            const REAL* oa = r->obst_actual + ((size_t)n * NO + b) * 3;
            const int in_range = FN(obstacle_distance)(h, st, hz, oa) < vis;                      /* :631-643 */
            for (int q = 0; q < 3; ++q) o[k + 3 * b + q] = in_range ? oa[q] : (REAL)c->track.obstacles[b][q];
            o[k + 3 * NO + b] = (REAL)in_range;
        }
        k += 4 * NO;
        o[k++] = (REAL)r->current_gate[(size_t)n * M + i];                                        /* :651 */
        if (c->race_mode == ADRP_RACE_COMPETE)                                                    /* :654-659 */
            for (int j = 0; j < M; ++j) {
                if (j == i) continue;
                const REAL* sj = h->state + ((size_t)n * M + j) * 20;
                REAL rpy[3];
                FN(euler)(sj + 3, rpy);
                for (int q = 0; q < 3; ++q) { o[k + q] = sj[q]; o[k + 3 + q] = rpy[q]; }
                k += 6;
            }
    }
}

/* MultiRaceAviary._collision (:552-562): any collision object within Bullet's contact margin */
static int FN(race_crashed)(FN(orc)* h, int n, int i)
{
    FN(orc_race)* r = h->race;
    const adrp_config* c = &h->cfg;
    const int M = h->M;
    const REAL margin = (REAL)c->track.contact_margin;
    const REAL* st = h->state + ((size_t)n * M + i) * 20;
    const REAL hz = FN(drone_hz)(h, st);
    for (int g = 0; g < r->G; ++g)
        if (FN(gate_distance)(h, st, hz, r->gates_actual + ((size_t)n * r->G + g) * 4, c->track.gate_type[g]) < margin) return 1;
    for (int b = 0; b < r->O; ++b)
        if (FN(obstacle_distance)(h, st, hz, r->obst_actual + ((size_t)n * r->O + b) * 3) < margin) return 1;
    if (st[2] - hz < margin) return 1;                                /* plane */
    if (c->race_mode == ADRP_RACE_COMPETE)                            /* :153-154 */
        for (int j = 0; j < M; ++j) {
            if (j == i) continue;
            const REAL* sj = h->state + ((size_t)n * M + j) * 20;
            REAL dxy = FN(gap_circle_circle)(st[0], st[1], h->D.coll_r, sj[0], sj[1], h->D.coll_r);
            if (FN(combine)(dxy, FABS(st[2] - sj[2]) - hz - FN(drone_hz)(h, sj)) < margin) return 1;
        }
    return 0;
}

static void FN(race_post_env)(FN(orc)* h, int n, REAL* obs, REAL* reward, uint8_t* te, uint8_t* tr, uint8_t* fl);

/* MultiRaceAviary.step (:171-270) for env n.
 * noise: NULL or (N, M, S, 7) = [wind force xyz (N), motor thrust noise x4 (N)] per substep. */
static void FN(race_step_env)(FN(orc)* h, int n, const REAL* action, const REAL* noise, REAL* obs, REAL* reward,
                              uint8_t* te, uint8_t* tr, uint8_t* fl)
{
    FN(orc_race)* r = h->race;
    const adrp_config* c = &h->cfg;
    const int M = h->M, S = h->S, phy = c->physics;
    const FN(orc_drone)* D = &h->D;
    /* :190-202 ndarray action -> FULLSTATE(pos, 0, 0, yaw, 0, step_counter); eliminated drones get STOP */
    for (int i = 0; i < M; ++i) {
        FN(mel_state)* m = r->mel + (size_t)n * M + i;
        if (r->latch[(size_t)n * M + i] & ADRP_FLAG_ELIMINATED) continue;
        const REAL* a = action + ((size_t)n * M + i) * h->A;
        if (c->action_type == ADRP_ACT_MEL_CMD) {
            /* (Command, args) tuples (MultiRaceAviary.py:197-205 -> low_level_control, MellingerControl.py:32-57): the send*
             * call queues the command and process_command_queue(args[-1]) (:292-303) pops it at once -- Stop(), then
             * UpdateTime(args[-1]) (the LAST argument, whatever it means for that command), then the command itself */
            const int cmd = (int)a[0];
            const REAL* g = a + 1;
            if (cmd == ADRP_CMD_NONE) continue;                      /* `else: continue`: no queue processing */
            static const int last_arg[] = {0, 13, 1, 2, 2, 1, 2, 2, 0, 5, 0};   /* index of args[-1] per command */
            FN(hl_stop)(&m->plan);
            FN(hl_update_time)(&m->plan, (float)g[last_arg[cmd]]);
            switch (cmd) {
            case ADRP_CMD_TAKEOFF: FN(hl_takeoff)(&m->plan, (float)g[0], (float)g[1], 1, 0.0f); m->hl_active = 1; continue;
            case ADRP_CMD_TAKEOFFYAW: FN(hl_takeoff)(&m->plan, (float)g[0], (float)g[1], 0, (float)g[2]); m->hl_active = 1; continue;
            case ADRP_CMD_TAKEOFFVEL: FN(hl_takeoff_vel)(&m->plan, (float)g[0], (float)g[1], g[2] != 0); m->hl_active = 1; continue;
            case ADRP_CMD_LAND: FN(hl_land)(&m->plan, (float)g[0], (float)g[1], 1, 0.0f); m->hl_active = 1; continue;
            case ADRP_CMD_LANDYAW: FN(hl_land)(&m->plan, (float)g[0], (float)g[1], 0, (float)g[2]); m->hl_active = 1; continue;
            case ADRP_CMD_LANDVEL: FN(hl_land_vel)(&m->plan, (float)g[0], (float)g[1], g[2] != 0); m->hl_active = 1; continue;
            case ADRP_CMD_STOP: FN(hl_stop)(&m->plan); m->hl_active = 1; continue;
            case ADRP_CMD_GOTO: FN(hl_go_to)(&m->plan, (float)g[0], (float)g[1], (float)g[2], (float)g[3], (float)g[4], g[5] != 0); m->hl_active = 1; continue;
            case ADRP_CMD_NOTIFY:                                    /* _notifySetpointStop (:693-699): TellState(self.state), */
                FN(hl_tell_state)(&m->plan, m->st_pos, m->st_vel, m->st_yaw_deg);   /* the state the last tick stored */
                m->hl_active = 1;
                continue;
            default: break;                                          /* ADRP_CMD_FULLSTATE: the 14 arguments follow */
            }
            a = g;
            m->hl_active = 0;                                        /* full_state_cmd_override = True (:543) */
        }
        const int full = c->action_type == ADRP_ACT_MEL_FULL || c->action_type == ADRP_ACT_MEL_CMD;   /* (Command.FULLSTATE, (pos, vel, acc, yaw, rpy_rate, t)) */
        for (int k = 0; k < 3; ++k) { m->sp_pos[k] = (float)a[k]; m->sp_vel[k] = 0; m->sp_acc[k] = 0; m->sp_rate_deg[k] = 0; }
        if (full)                                                    /* _sendFullStateCmd (MellingerControl.py:510-527) */
            for (int k = 0; k < 3; ++k) {
                m->sp_vel[k] = (float)a[3 + k]; m->sp_acc[k] = (float)a[6 + k];
                m->sp_rate_deg[k] = (float)(a[10 + k] * (REAL)(180.0 / 3.14159265358979323846));   /* RAD_TO_DEG */
            }
        REAL yaw = (c->wrappers & ADRP_WRAP_OBS) ? (REAL)0 : a[full ? 9 : 3];   /* wrapper.py:52-57; get_quaternion_from_euler(0, 0, yaw) (:525) */
        m->sp_quat[0] = (float)(SIN((REAL)0) * COS((REAL)0) * COS(yaw / 2) - COS((REAL)0) * SIN((REAL)0) * SIN(yaw / 2));
        m->sp_quat[1] = (float)(COS((REAL)0) * SIN((REAL)0) * COS(yaw / 2) + SIN((REAL)0) * COS((REAL)0) * SIN(yaw / 2));
        m->sp_quat[2] = (float)(COS((REAL)0) * COS((REAL)0) * SIN(yaw / 2) - SIN((REAL)0) * SIN((REAL)0) * COS(yaw / 2));
        m->sp_quat[3] = (float)(COS((REAL)0) * COS((REAL)0) * COS(yaw / 2) + SIN((REAL)0) * SIN((REAL)0) * SIN(yaw / 2));
        m->sp_yaw_abs = 0;                                           /* mode.quat = modeAbs, mode.yaw = modeDisable (:536-540) */
    }
    /* RewardWrapper.reset (wrapper.py:96-101) reads the reset observation of drone 0: current gate id, the
     * observed pose of gate 0 and the position.  Evaluated here, on the state the episode starts from. */
    if ((c->wrappers & ADRP_WRAP_REWARD) && r->shaping[7 * (size_t)n] < 0) {
        REAL* sh = r->shaping + 7 * (size_t)n;
        const REAL* st = h->state + (size_t)n * M * 20;
        sh[0] = (REAL)r->current_gate[(size_t)n * M];
        for (int k = 0; k < 3; ++k) sh[1 + k] = 0;
        if (r->G > 0) {
            const REAL* ga = r->gates_actual + (size_t)n * r->G * 4;
            const int in_range = FN(gate_distance)(h, st, FN(drone_hz)(h, st), ga, c->track.gate_type[0]) < (REAL)c->track.visibility_range;
            for (int k = 0; k < 3; ++k) sh[1 + k] = in_range ? ga[k] : (REAL)c->track.gates[0][k];
        }
        for (int k = 0; k < 3; ++k) sh[4 + k] = st[k];
    }
    for (int s = 0; s < S; ++s) {
        REAL* env = h->state + (size_t)n * M * 20;
        REAL cache[ADRP_MAX_DRONES * 20];
        memcpy(cache, env, sizeof(REAL) * 20 * M);
        /* per-substep disturbance draws: injected, or Philox(seed; env, drone, tick, episode) */
        REAL nz[ADRP_MAX_DRONES][7];
        for (int i = 0; i < M; ++i) {
            for (int k = 0; k < 7; ++k) nz[i][k] = 0;
            if (noise) {
                for (int k = 0; k < 7; ++k) nz[i][k] = noise[(((size_t)n * M + i) * S + s) * 7 + k];
            } else if (c->rnd.disturbances) {
                uint32_t o0[4], o1[4];
                const uint32_t tick = (uint32_t)(h->step_counter[n] + s), ep = (uint32_t)r->episode[n];
                FN(philox)((uint32_t)(c->env_offset + (uint64_t)n), (uint32_t)(64 + i), tick, ep, c->seed, o0);
                FN(philox)((uint32_t)(c->env_offset + (uint64_t)n), (uint32_t)(128 + i), tick, ep, c->seed, o1);
                for (int k = 0; k < 3; ++k)
                    nz[i][k] = (REAL)c->rnd.wind_lo[k] + FN(u01)(o0[k]) * ((REAL)c->rnd.wind_hi[k] - (REAL)c->rnd.wind_lo[k]);
                /* Box-Muller: two pairs -> four normals */
                REAL r0 = SQRT((REAL)-2 * LOG(FN(u01_open)(o1[0]))), r1 = SQRT((REAL)-2 * LOG(FN(u01_open)(o1[2])));
                REAL a0 = (REAL)6.283185307179586 * FN(u01)(o1[1]), a1 = (REAL)6.283185307179586 * FN(u01)(o1[3]);
                nz[i][3] = (REAL)c->rnd.action_noise_std * r0 * COS(a0);
                nz[i][4] = (REAL)c->rnd.action_noise_std * r0 * SIN(a0);
                nz[i][5] = (REAL)c->rnd.action_noise_std * r1 * COS(a1);
                nz[i][6] = (REAL)c->rnd.action_noise_std * r1 * SIN(a1);
            }
        }
        /* _apply_physics(self.rpms, self.prev_rpms) (:510-548) */
        for (int i = 0; i < M; ++i) {
            REAL* st = env + 20 * i;
            FN(mel_state)* m = r->mel + (size_t)n * M + i;
            REAL R[3][3], gnd[4] = {0, 0, 0, 0}, drag_w[3] = {0, 0, 0}, dw = 0;
            FN(rotmat)(cache + 20 * i + 3, R);
            if (phy & ADRP_PHY_GND) FN(ground_effect)(D, cache + 20 * i, R, m->rpms, gnd);
            if (phy & ADRP_PHY_DRAG) FN(drag)(D, cache + 20 * i, cache + 20 * i + 16, drag_w);    /* prev_rpms */
            if (phy & ADRP_PHY_DW) dw = FN(downwash)(D, cache, M, i);
            if (phy & ADRP_PHY_DYN) {
                FN(dynamics)(D, c->drone_model, st, m->rpms, gnd, drag_w, dw);
            } else {
                FN(orc_drone) Db = *D;                               /* per-drone Bullet body (changeDynamics, :426-431) */
                const REAL* body = r->body + ((size_t)n * M + i) * 4;
                Db.pyb_mass = body[0]; Db.pyb_J[0] = body[1]; Db.pyb_J[1] = body[2]; Db.pyb_J[2] = body[3];
                FN(pyb_step)(&Db, phy, st, m->rpms, gnd, drag_w, dw, nz[i]);
            }
        }
        /* controller round trip (:231-254) */
        for (int i = 0; i < M; ++i) {
            REAL* st = env + 20 * i;
            FN(mel_state)* m = r->mel + (size_t)n * M + i;
            if (r->latch[(size_t)n * M + i] & ADRP_FLAG_ELIMINATED) {
                for (int k = 0; k < 4; ++k) { st[16 + k] = 0; m->rpms[k] = 0; }   /* :233-235 */
                continue;
            }
            for (int k = 0; k < 4; ++k) st[16 + k] = m->rpms[k];     /* prev_rpms[i] = rpms[i] (:251) */
            FN(mellinger_tick)(h, m, st, nz[i] + 3);                 /* rpms[i] = recv() (:254) */
        }
    }
    FN(race_post_env)(h, n, obs, reward, te, tr, fl);
}

/* the part of MultiRaceAviary.step after the substep loop (:257-268): gate progress, observation, flags */
static void FN(race_post_env)(FN(orc)* h, int n, REAL* obs, REAL* reward, uint8_t* te, uint8_t* tr, uint8_t* fl)
{
    FN(orc_race)* r = h->race;
    const adrp_config* c = &h->cfg;
    const int M = h->M, S = h->S, phy = c->physics;
    FN(gate_progress)(h, n);                                         /* :257-258 */
    FN(race_write_obs)(h, n, obs);                                   /* :261 */
    /* _computeTerminated (:674-698) */
    int all_done = 1;
    for (int i = 0; i < M; ++i) {
        const REAL* st = h->state + ((size_t)n * M + i) * 20;
        uint8_t* L = r->latch + (size_t)n * M + i;
        uint8_t f = 0;
        for (int k = 0; k < 3; ++k) {
            if (FABS(st[k]) > (REAL)c->track.bounds_hi[k]) f |= ADRP_FLAG_OUT_OF_BOUNDS;   /* :684 */
            if (FABS(st[10 + k]) > (REAL)20) f |= ADRP_FLAG_UNSTABLE;                       /* :685 */
        }
        if (!(phy & ADRP_PHY_DYN) && FN(race_crashed)(h, n, i)) f |= ADRP_FLAG_CRASHED;     /* :686, E9 */
        if (f) *L |= ADRP_FLAG_ELIMINATED;
        if (r->mel[(size_t)n * M + i].error) f |= ADRP_FLAG_TUMBLE_STOP;
        if (fl) fl[(size_t)n * M + i] = (uint8_t)(f | *L);
        if (!(*L & (ADRP_FLAG_ELIMINATED | ADRP_FLAG_FINISHED))) all_done = 0;
    }
    reward[n] = 0;                                                   /* :665-670 */
    if (c->wrappers & ADRP_WRAP_REWARD) {
        /* RewardWrapper._compute_reward (wrapper.py:121-186) on drone 0's row of this step's observation.
         * Declared where the reference raises: info has no "task_completed" key in this fork (KeyError once the env
         * terminates) -> drone 0's FINISHED latch; gate id == number of gates has no entry in its pose table
         * (KeyError) -> the target is kept. */
        REAL* sh = r->shaping + 7 * (size_t)n;
        const REAL* o = obs + (size_t)n * M * h->O;
        const int gate_id = (int)r->current_gate[(size_t)n * M];
        REAL r_passed = 0, r_collision = 0, r_lab = 0;
        if (gate_id > ((int)sh[0]) % 4) {
            sh[0] = (REAL)gate_id;
            if (gate_id < r->G) for (int k = 0; k < 3; ++k) sh[1 + k] = o[12 + 4 * gate_id + k];
            r_passed = 5;
        }
        const int completed = (r->latch[(size_t)n * M] & ADRP_FLAG_FINISHED) != 0;
        if (all_done && !completed) r_collision = -1;
        if (all_done && completed) r_lab = 10;
        const REAL dpx = sh[1] - sh[4], dpy = sh[2] - sh[5], dcx = sh[1] - o[0], dcy = sh[2] - o[1];
        const REAL prog_xy = SQRT(dpx * dpx + dpy * dpy) - SQRT(dcx * dcx + dcy * dcy);
        const REAL prog_z = FABS(sh[3] - sh[6]) - FABS(sh[3] - o[2]);
        reward[n] = prog_xy + prog_z + r_passed + r_collision + r_lab;
        for (int k = 0; k < 3; ++k) sh[4 + k] = o[k];
    }
    if ((c->wrappers & ADRP_WRAP_OBS) && r->current_gate[(size_t)n * M] >= 2) all_done = 1;   /* wrapper.py:60-61 */
    te[n] = (uint8_t)all_done;
    tr[n] = (uint8_t)((double)h->step_counter[n] / (double)c->pyb_freq > c->episode_len_sec);   /* :709 */
    h->step_counter[n] += S;
    if (c->auto_reset && (te[n] || tr[n])) { FN(init_env)(h, n); FN(race_init_env)(h, n, NULL); }
}
