/* TEST INFRASTRUCTURE -- DECLARED restatement of the firmware's high-level commander (pycffirmware, un-vendored:
 * crazyflie-firmware src/modules/src/crtp_commander_high_level.c, planner.c, pptraj.c, math3d.h), the part of it
 * control/MellingerControl.py reaches through crtpCommanderHighLevel* (:139,:150,:292-303,:369-374,:547-699).
 * PARITY UNPINNED: restated from the published algorithm, no reference test or vector pins a value; the golden run
 * tests/golden/race_commander_*.npz pins the REFERENCE-OWNED logic around it (command queue, Stop() + UpdateTime(args[-1])
 * before every command, full_state_cmd_override switching) on top of this model.  float32 like the firmware; libm calls
 * through FW_SINF / FW_COSF (adrp_oracle.c) so that the CUDA parity build can reproduce them bit for bit.
 *
 * What is modelled: one-piece 7th-order "no jerk" polynomials (piecewise_plan_7th_order_no_jerk) for take-off / landing /
 * go-to, the planner's state machine (IDLE / FLYING / LANDING; plan_takeoff refuses unless IDLE, plan_land refuses when
 * IDLE or LANDING, go-to is allowed in every state), evaluation with the differential-flatness body rates (poly4d_eval),
 * and the commander's last-setpoint memory (pos, vel, yaw).  Not modelled (unreachable through MellingerControl.py):
 * uploaded / compressed trajectories, reversed playback, group masks.  Included by adrp_oracle_race.h. */

enum { FN(PLAN_IDLE) = 0, FN(PLAN_FLYING) = 1, FN(PLAN_LANDING) = 2 };

typedef struct FN(planner) {
    int32_t state;                 /* enum trajectory_state */
    float p[4][8];                 /* struct poly4d: x, y, z, yaw coefficients, lowest order first (PP_SIZE = 8) */
    float duration;                /* of the one piece */
    float t_begin;                 /* piecewise_traj.t_begin */
    float pos[3], vel[3], yaw;     /* crtp_commander_high_level.c statics: last known setpoint (or state estimate) */
    float t;                       /* commander time as set by crtpCommanderHighLevelUpdateTime (pycffirmware's clock) */
} FN(planner);

typedef struct FN(traj_eval) { float pos[3], vel[3], acc[3], omega[3], yaw; int valid; } FN(traj_eval);

/* pptraj.c poly7_nojerk */
static void FN(poly7_nojerk)(float poly[8], float T, float x0, float dx0, float ddx0, float xf, float dxf, float ddxf)
{
    if (T <= 0.0f) {
        poly[0] = xf; poly[1] = dxf; poly[2] = ddxf / 2;
        for (int i = 3; i < 8; ++i) poly[i] = 0;
        return;
    }
    float T2 = T * T, T3 = T2 * T, T4 = T3 * T, T5 = T4 * T, T6 = T5 * T, T7 = T6 * T;
    poly[0] = x0;
    poly[1] = dx0;
    poly[2] = ddx0 / 2;
    poly[3] = 0;
    poly[4] = -(5 * (14 * x0 - 14 * xf + 8 * T * dx0 + 6 * T * dxf + 2 * T2 * ddx0 - T2 * ddxf)) / (2 * T4);
    poly[5] = (84 * x0 - 84 * xf + 45 * T * dx0 + 39 * T * dxf + 10 * T2 * ddx0 - 7 * T2 * ddxf) / T5;
    poly[6] = -(140 * x0 - 140 * xf + 72 * T * dx0 + 68 * T * dxf + 15 * T2 * ddx0 - 13 * T2 * ddxf) / (2 * T6);
    poly[7] = (2 * (10 * x0 - 10 * xf + 5 * T * dx0 + 5 * T * dxf + ddx0 * T2 - ddxf * T2)) / T7;
}

/* pptraj.c polyval (Horner) / polyder */
static float FN(polyval)(const float p[8], float t)
{
    float x = 0.0f;
    for (int i = 7; i >= 0; --i) x = x * t + p[i];
    return x;
}
static void FN(polyder)(float p[8])
{
    for (int i = 1; i <= 7; ++i) p[i - 1] = i * p[i];
    p[7] = 0;
}

/* pptraj.c poly4d_eval: flat outputs, their derivatives and the body rates from jerk (differential flatness) */
static FN(traj_eval) FN(poly4d_eval)(const float pin[4][8], float t)
{
    FN(traj_eval) out;
    float d[4][8];
    memcpy(d, pin, sizeof(d));
    for (int k = 0; k < 3; ++k) out.pos[k] = FN(polyval)(d[k], t);
    out.yaw = FN(polyval)(d[3], t);
    for (int k = 0; k < 4; ++k) FN(polyder)(d[k]);
    for (int k = 0; k < 3; ++k) out.vel[k] = FN(polyval)(d[k], t);
    float dyaw = FN(polyval)(d[3], t);
    for (int k = 0; k < 4; ++k) FN(polyder)(d[k]);
    for (int k = 0; k < 3; ++k) out.acc[k] = FN(polyval)(d[k], t);
    for (int k = 0; k < 4; ++k) FN(polyder)(d[k]);
    float jerk[3];
    for (int k = 0; k < 3; ++k) jerk[k] = FN(polyval)(d[k], t);
    float thrust[3] = {out.acc[0], out.acc[1], out.acc[2] + 9.81f};                 /* GRAV */
    float tm = sqrtf(thrust[0] * thrust[0] + thrust[1] * thrust[1] + thrust[2] * thrust[2]);
    float zb[3] = {thrust[0] / tm, thrust[1] / tm, thrust[2] / tm};                 /* vnormalize */
    float xw[3] = {FW_COSF(out.yaw), FW_SINF(out.yaw), 0};
    float yb[3] = {zb[1] * xw[2] - zb[2] * xw[1], zb[2] * xw[0] - zb[0] * xw[2], zb[0] * xw[1] - zb[1] * xw[0]};
    float ym = sqrtf(yb[0] * yb[0] + yb[1] * yb[1] + yb[2] * yb[2]);
    yb[0] /= ym; yb[1] /= ym; yb[2] /= ym;
    float xb[3] = {yb[1] * zb[2] - yb[2] * zb[1], yb[2] * zb[0] - yb[0] * zb[2], yb[0] * zb[1] - yb[1] * zb[0]};
    float jz = jerk[0] * zb[0] + jerk[1] * zb[1] + jerk[2] * zb[2];                 /* vorthunit(jerk, z_body) */
    float jo[3] = {jerk[0] - jz * zb[0], jerk[1] - jz * zb[1], jerk[2] - jz * zb[2]};
    float inv = 1.0f / tm;
    float hw[3] = {inv * jo[0], inv * jo[1], inv * jo[2]};
    out.omega[0] = -(hw[0] * yb[0] + hw[1] * yb[1] + hw[2] * yb[2]);
    out.omega[1] = hw[0] * xb[0] + hw[1] * xb[1] + hw[2] * xb[2];
    out.omega[2] = zb[2] * dyaw;
    out.valid = 1;
    return out;
}

/* pptraj.c piecewise_eval for the planner's one-piece trajectory (timescale 1, shift 0) */
static FN(traj_eval) FN(piecewise_eval)(const FN(planner)* p, float t)
{
    t = t - p->t_begin;
    if (t <= p->duration * 1.0f) return FN(poly4d_eval)(p->p, t);           /* also for t < 0: the polynomial is extrapolated */
    FN(traj_eval) ev = FN(poly4d_eval)(p->p, p->duration);                   /* the trajectory has ended */
    for (int k = 0; k < 3; ++k) { ev.vel[k] = 0; ev.acc[k] = 0; ev.omega[k] = 0; }
    return ev;
}

/* planner.c plan_current_goal: LANDING turns IDLE once finished and still falls through into this call's evaluation */
static FN(traj_eval) FN(plan_current_goal)(FN(planner)* p, float t)
{
    FN(traj_eval) ev;
    memset(&ev, 0, sizeof(ev));                                               /* traj_eval_invalid() */
    if (p->state == FN(PLAN_IDLE)) return ev;
    if (p->state == FN(PLAN_LANDING) && (t - p->t_begin) >= p->duration) p->state = FN(PLAN_IDLE);   /* plan_is_finished */
    return FN(piecewise_eval)(p, t);
}

/* math3d.h normalize_radians / shortest_signed_angle_radians */
static float FN(normalize_radians)(float r)
{
    const float PI_F = 3.14159265358979323846f;
    float signed_pi = copysignf(PI_F, r);
    return fmodf(r + signed_pi, 2 * PI_F) - signed_pi;
}
static float FN(shortest_signed_angle)(float start, float goal)
{
    const float PI_F = 3.14159265358979323846f;
    float diff = goal - start, x = diff + PI_F, n = 2 * PI_F;
    return (x - floorf(x / n) * n) - PI_F;                                   /* fmodf_floored */
}

/* pptraj.c piecewise_plan_7th_order_no_jerk + the bookkeeping every plan_* does afterwards */
static void FN(plan_piece)(FN(planner)* p, float duration, const float p0[3], float y0, const float v0[3], float dy0,
                           const float a0[3], const float p1[3], float y1, int state, float t)
{
    p->duration = duration;
    for (int k = 0; k < 3; ++k) FN(poly7_nojerk)(p->p[k], duration, p0[k], v0[k], a0[k], p1[k], 0, 0);
    FN(poly7_nojerk)(p->p[3], duration, y0, dy0, 0, y1, 0, 0);
    p->state = state;
    p->t_begin = t;
}

/* planner.c plan_takeoff / plan_land (plan_takeoff_or_landing) */
static int FN(plan_takeoff_or_land)(FN(planner)* p, int landing, float hover_height, float hover_yaw, float duration)
{
    if (!landing && p->state != FN(PLAN_IDLE)) return 1;
    if (landing && (p->state == FN(PLAN_IDLE) || p->state == FN(PLAN_LANDING))) return 1;
    const float zero[3] = {0, 0, 0};
    float hover_pos[3] = {p->pos[0], p->pos[1], hover_height};
    hover_yaw = FN(normalize_radians)(hover_yaw);
    float curr_yaw = FN(normalize_radians)(p->yaw);
    float goal_yaw = curr_yaw + FN(shortest_signed_angle)(curr_yaw, hover_yaw);
    FN(plan_piece)(p, duration, p->pos, curr_yaw, zero, 0, zero, hover_pos, goal_yaw,
                   landing ? FN(PLAN_LANDING) : FN(PLAN_FLYING), p->t);
    return 0;
}

/* crtp_commander_high_level.c */
static void FN(hl_init)(FN(planner)* p) { memset(p, 0, sizeof(*p)); }             /* plan_init, pos = vel = 0, yaw = 0 */
static void FN(hl_stop)(FN(planner)* p) { p->state = FN(PLAN_IDLE); }              /* plan_stop */
static void FN(hl_update_time)(FN(planner)* p, float t) { p->t = t; }
static void FN(hl_tell_state)(FN(planner)* p, const float pos[3], const float vel[3], float yaw_deg)
{
    for (int k = 0; k < 3; ++k) { p->pos[k] = pos[k]; p->vel[k] = vel[k]; }
    p->yaw = (3.14159265358979323846f / 180.0f) * yaw_deg;                        /* radians(state->attitude.yaw) */
}
/* takeoff2 / land2: use_current_yaw for the plain commands, the given yaw for the *Yaw variants */
static int FN(hl_takeoff)(FN(planner)* p, float height, float duration, int use_current_yaw, float yaw)
{
    return FN(plan_takeoff_or_land)(p, 0, height, use_current_yaw ? p->yaw : yaw, duration);
}
static int FN(hl_land)(FN(planner)* p, float height, float duration, int use_current_yaw, float yaw)
{
    return FN(plan_takeoff_or_land)(p, 1, height, use_current_yaw ? p->yaw : yaw, duration);
}
/* takeoff_with_velocity / land_with_velocity (defaultTakeoffVelocity = defaultLandingVelocity = 0.5) */
static int FN(hl_takeoff_vel)(FN(planner)* p, float height, float velocity, int relative)
{
    if (relative) height += p->pos[2];
    velocity = velocity > 0 ? velocity : 0.5f;
    return FN(plan_takeoff_or_land)(p, 0, height, p->yaw, fabsf(height - p->pos[2]) / velocity);
}
static int FN(hl_land_vel)(FN(planner)* p, float height, float velocity, int relative)
{
    if (relative) height = p->pos[2] - height;
    velocity = velocity > 0 ? velocity : 0.5f;
    return FN(plan_takeoff_or_land)(p, 1, height, p->yaw, fabsf(height - p->pos[2]) / velocity);
}
/* go_to: from the last setpoint memory when the planner is stopped, else from the current goal (plan_go_to) */
static int FN(hl_go_to)(FN(planner)* p, float x, float y, float z, float yaw, float duration, int relative)
{
    FN(traj_eval) cur;
    memset(&cur, 0, sizeof(cur));
    if (p->state == FN(PLAN_IDLE)) {
        for (int k = 0; k < 3; ++k) { cur.pos[k] = p->pos[k]; cur.vel[k] = p->vel[k]; }
        cur.yaw = p->yaw;
    } else {
        cur = FN(plan_current_goal)(p, p->t);
    }
    float hover_pos[3] = {x, y, z}, hover_yaw = yaw;
    if (relative) {
        for (int k = 0; k < 3; ++k) hover_pos[k] += cur.pos[k];
        hover_yaw += cur.yaw;
    }
    float curr_yaw = FN(normalize_radians)(cur.yaw);
    hover_yaw = FN(normalize_radians)(hover_yaw);
    float goal_yaw = curr_yaw + FN(shortest_signed_angle)(curr_yaw, hover_yaw);
    FN(plan_piece)(p, duration, cur.pos, curr_yaw, cur.vel, cur.omega[2], cur.acc, hover_pos, goal_yaw, FN(PLAN_FLYING), p->t);
    return 0;
}

/* crtpCommanderHighLevelGetSetpoint.  out[14] = pos3 vel3 acc3 rate_deg3 yaw_deg, valid flag; state_* = the estimate
 * handed in (state->position, ->velocity, ->attitude.yaw in degrees).  Returns 1 when the setpoint was written. */
static int FN(hl_get_setpoint)(FN(planner)* p, const float state_pos[3], const float state_vel[3], float state_yaw_deg, float out[13])
{
    FN(traj_eval) ev = FN(plan_current_goal)(p, p->t);
    if (p->state == FN(PLAN_IDLE)) FN(hl_tell_state)(p, state_pos, state_vel, state_yaw_deg);   /* "on the ground": track the estimate */
    if (!ev.valid) return 0;
    const float DEG = 180.0f / 3.14159265358979323846f;
    for (int k = 0; k < 3; ++k) { out[k] = ev.pos[k]; out[3 + k] = ev.vel[k]; out[6 + k] = ev.acc[k]; out[9 + k] = DEG * ev.omega[k]; }
    out[12] = DEG * ev.yaw;
    for (int k = 0; k < 3; ++k) { p->pos[k] = ev.pos[k]; p->vel[k] = ev.vel[k]; }                /* store the last setpoint */
    p->yaw = ev.yaw;
    return 1;
}
