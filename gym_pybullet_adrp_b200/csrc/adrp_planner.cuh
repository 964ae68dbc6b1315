// adrp_planner.cuh -- the firmware's high-level commander as device functions (ADRP_ACT_MEL_CMD): what
// control/MellingerControl.py reaches through crtpCommanderHighLevel* (:139,:150,:292-303,:369-374,:547-699) when a
// controller sends Command.TAKEOFF / LAND / GOTO / STOP / NOTIFY tuples (user_controller/HardCodedController.py:158-188).
//
// DECLARED restatement of un-vendored pycffirmware code (crtp_commander_high_level.c, planner.c, pptraj.c, math3d.h),
// parity unpinned like fw_mellinger: one-piece 7th-order "no jerk" polynomials, the planner's IDLE / FLYING / LANDING state
// machine, evaluation with differential-flatness body rates, the commander's last-setpoint memory.  float32 in both
// builds, statement order fixed so that the CPU checker can match it; the fp64 parity translation unit is compiled with
// -fmad=false and calls sin / cos through the double-precision functions (FwMath<double>).
#pragma once
#include "adrp_mellinger.cuh"

namespace adrp {

enum { PLAN_IDLE = 0, PLAN_FLYING = 1, PLAN_LANDING = 2 };

struct Planner {
    int state;                     // enum trajectory_state
    float p[4][8];                 // struct poly4d: x, y, z, yaw coefficients, lowest order first
    float duration, t_begin;       // of the one piece
    float pos[3], vel[3], yaw;     // commander statics: last known setpoint (or state estimate)
    float t;                       // commander time (crtpCommanderHighLevelUpdateTime)
};

struct TrajEval { float pos[3], vel[3], acc[3], omega[3], yaw; int valid; };

// Per-drone high-level state of the Mellinger wrapper: planner, full_state_cmd_override (inverted), the setpoint struct
// (which persists from one control step to the next: Command.NONE sends nothing) and state_t as the last tick stored it.
struct HlState {
    Planner plan;
    int active;                    // not full_state_cmd_override
    MelSetpoint sp;
    float st_pos[3], st_vel[3], st_yaw_deg;
};

__device__ inline MelSetpoint& hl_setpoint(HlState& h) { return h.sp; }
constexpr int kHlWords = (int)(sizeof(HlState) / 4);     // every member is a 4-byte int or float: stored as SoA words in HBM
static_assert(sizeof(HlState) % 4 == 0, "HlState is stored word by word");

// pptraj.c poly7_nojerk
__device__ inline void poly7_nojerk(float poly[8], float T, float x0, float dx0, float ddx0, float xf, float dxf, float ddxf)
{
    if (T <= 0.0f) {
        poly[0] = xf; poly[1] = dxf; poly[2] = ddxf / 2;
        for (int i = 3; i < 8; ++i) poly[i] = 0;
        return;
    }
    const float T2 = T * T, T3 = T2 * T, T4 = T3 * T, T5 = T4 * T, T6 = T5 * T, T7 = T6 * T;
    poly[0] = x0;
    poly[1] = dx0;
    poly[2] = ddx0 / 2;
    poly[3] = 0;
    poly[4] = -(5 * (14 * x0 - 14 * xf + 8 * T * dx0 + 6 * T * dxf + 2 * T2 * ddx0 - T2 * ddxf)) / (2 * T4);
    poly[5] = (84 * x0 - 84 * xf + 45 * T * dx0 + 39 * T * dxf + 10 * T2 * ddx0 - 7 * T2 * ddxf) / T5;
    poly[6] = -(140 * x0 - 140 * xf + 72 * T * dx0 + 68 * T * dxf + 15 * T2 * ddx0 - 13 * T2 * ddxf) / (2 * T6);
    poly[7] = (2 * (10 * x0 - 10 * xf + 5 * T * dx0 + 5 * T * dxf + ddx0 * T2 - ddxf * T2)) / T7;
}

__device__ inline float plan_polyval(const float p[8], float t)
{
    float x = 0.0f;
    for (int i = 7; i >= 0; --i) x = x * t + p[i];
    return x;
}
__device__ inline void plan_polyder(float p[8])
{
    for (int i = 1; i <= 7; ++i) p[i - 1] = i * p[i];
    p[7] = 0;
}

// pptraj.c poly4d_eval
template <typename T>
__device__ inline TrajEval poly4d_eval(const float pin[4][8], float t)
{
    TrajEval out;
    float d[4][8];
    for (int a = 0; a < 4; ++a)
        for (int i = 0; i < 8; ++i) d[a][i] = pin[a][i];
    for (int k = 0; k < 3; ++k) out.pos[k] = plan_polyval(d[k], t);
    out.yaw = plan_polyval(d[3], t);
    for (int k = 0; k < 4; ++k) plan_polyder(d[k]);
    for (int k = 0; k < 3; ++k) out.vel[k] = plan_polyval(d[k], t);
    const float dyaw = plan_polyval(d[3], t);
    for (int k = 0; k < 4; ++k) plan_polyder(d[k]);
    for (int k = 0; k < 3; ++k) out.acc[k] = plan_polyval(d[k], t);
    for (int k = 0; k < 4; ++k) plan_polyder(d[k]);
    float jerk[3];
    for (int k = 0; k < 3; ++k) jerk[k] = plan_polyval(d[k], t);
    const float thrust[3] = {out.acc[0], out.acc[1], out.acc[2] + 9.81f};
    const float tm = sqrtf(thrust[0] * thrust[0] + thrust[1] * thrust[1] + thrust[2] * thrust[2]);
    const float zb[3] = {thrust[0] / tm, thrust[1] / tm, thrust[2] / tm};
    float sy, cy;
    FwMath<T>::sincos(out.yaw, &sy, &cy);
    const float xw[3] = {cy, sy, 0};
    float yb[3] = {zb[1] * xw[2] - zb[2] * xw[1], zb[2] * xw[0] - zb[0] * xw[2], zb[0] * xw[1] - zb[1] * xw[0]};
    const float ym = sqrtf(yb[0] * yb[0] + yb[1] * yb[1] + yb[2] * yb[2]);
    yb[0] /= ym; yb[1] /= ym; yb[2] /= ym;
    const float xb[3] = {yb[1] * zb[2] - yb[2] * zb[1], yb[2] * zb[0] - yb[0] * zb[2], yb[0] * zb[1] - yb[1] * zb[0]};
    const float jz = jerk[0] * zb[0] + jerk[1] * zb[1] + jerk[2] * zb[2];
    const float jo[3] = {jerk[0] - jz * zb[0], jerk[1] - jz * zb[1], jerk[2] - jz * zb[2]};
    const float inv = 1.0f / tm;
    const float hw[3] = {inv * jo[0], inv * jo[1], inv * jo[2]};
    out.omega[0] = -(hw[0] * yb[0] + hw[1] * yb[1] + hw[2] * yb[2]);
    out.omega[1] = hw[0] * xb[0] + hw[1] * xb[1] + hw[2] * xb[2];
    out.omega[2] = zb[2] * dyaw;
    out.valid = 1;
    return out;
}

// pptraj.c piecewise_eval (one piece, timescale 1, shift 0)
template <typename T>
__device__ inline TrajEval piecewise_eval(const Planner& p, float t)
{
    t = t - p.t_begin;
    if (t <= p.duration * 1.0f) return poly4d_eval<T>(p.p, t);
    TrajEval ev = poly4d_eval<T>(p.p, p.duration);
    for (int k = 0; k < 3; ++k) { ev.vel[k] = 0; ev.acc[k] = 0; ev.omega[k] = 0; }
    return ev;
}

// planner.c plan_current_goal
template <typename T>
__device__ inline TrajEval plan_current_goal(Planner& p, float t)
{
    TrajEval ev = {};
    if (p.state == PLAN_IDLE) return ev;
    if (p.state == PLAN_LANDING && (t - p.t_begin) >= p.duration) p.state = PLAN_IDLE;
    return piecewise_eval<T>(p, t);
}

// math3d.h normalize_radians / shortest_signed_angle_radians
__device__ inline float plan_normalize_radians(float r)
{
    const float PI_F = 3.14159265358979323846f;
    const float signed_pi = copysignf(PI_F, r);
    return fmodf(r + signed_pi, 2 * PI_F) - signed_pi;
}
__device__ inline float plan_shortest_signed_angle(float start, float goal)
{
    const float PI_F = 3.14159265358979323846f;
    const float diff = goal - start, x = diff + PI_F, n = 2 * PI_F;
    return (x - floorf(x / n) * n) - PI_F;
}

// pptraj.c piecewise_plan_7th_order_no_jerk + the bookkeeping every plan_* does afterwards
__device__ inline void plan_piece(Planner& p, float duration, const float p0[3], float y0, const float v0[3], float dy0,
                                  const float a0[3], const float p1[3], float y1, int state, float t)
{
    p.duration = duration;
    for (int k = 0; k < 3; ++k) poly7_nojerk(p.p[k], duration, p0[k], v0[k], a0[k], p1[k], 0, 0);
    poly7_nojerk(p.p[3], duration, y0, dy0, 0, y1, 0, 0);
    p.state = state;
    p.t_begin = t;
}

// planner.c plan_takeoff / plan_land
__device__ inline int plan_takeoff_or_land(Planner& p, int landing, float hover_height, float hover_yaw, float duration)
{
    if (!landing && p.state != PLAN_IDLE) return 1;
    if (landing && (p.state == PLAN_IDLE || p.state == PLAN_LANDING)) return 1;
    const float zero[3] = {0, 0, 0};
    const float hover_pos[3] = {p.pos[0], p.pos[1], hover_height};
    hover_yaw = plan_normalize_radians(hover_yaw);
    const float curr_yaw = plan_normalize_radians(p.yaw);
    const float goal_yaw = curr_yaw + plan_shortest_signed_angle(curr_yaw, hover_yaw);
    const float start[3] = {p.pos[0], p.pos[1], p.pos[2]};
    plan_piece(p, duration, start, curr_yaw, zero, 0, zero, hover_pos, goal_yaw, landing ? PLAN_LANDING : PLAN_FLYING, p.t);
    return 0;
}

// crtp_commander_high_level.c
__device__ inline void hl_tell_state(Planner& p, const float pos[3], const float vel[3], float yaw_deg)
{
    for (int k = 0; k < 3; ++k) { p.pos[k] = pos[k]; p.vel[k] = vel[k]; }
    p.yaw = (3.14159265358979323846f / 180.0f) * yaw_deg;
}
__device__ inline int hl_takeoff_vel(Planner& p, float height, float velocity, int relative)
{
    if (relative) height += p.pos[2];
    velocity = velocity > 0 ? velocity : 0.5f;
    return plan_takeoff_or_land(p, 0, height, p.yaw, fabsf(height - p.pos[2]) / velocity);
}
__device__ inline int hl_land_vel(Planner& p, float height, float velocity, int relative)
{
    if (relative) height = p.pos[2] - height;
    velocity = velocity > 0 ? velocity : 0.5f;
    return plan_takeoff_or_land(p, 1, height, p.yaw, fabsf(height - p.pos[2]) / velocity);
}
template <typename T>
__device__ inline int hl_go_to(Planner& p, float x, float y, float z, float yaw, float duration, int relative)
{
    TrajEval cur = {};
    if (p.state == PLAN_IDLE) {
        for (int k = 0; k < 3; ++k) { cur.pos[k] = p.pos[k]; cur.vel[k] = p.vel[k]; }
        cur.yaw = p.yaw;
    } else {
        cur = plan_current_goal<T>(p, p.t);
    }
    float hover_pos[3] = {x, y, z}, hover_yaw = yaw;
    if (relative) {
        for (int k = 0; k < 3; ++k) hover_pos[k] += cur.pos[k];
        hover_yaw += cur.yaw;
    }
    const float curr_yaw = plan_normalize_radians(cur.yaw);
    hover_yaw = plan_normalize_radians(hover_yaw);
    const float goal_yaw = curr_yaw + plan_shortest_signed_angle(curr_yaw, hover_yaw);
    plan_piece(p, duration, cur.pos, curr_yaw, cur.vel, cur.omega[2], cur.acc, hover_pos, goal_yaw, PLAN_FLYING, p.t);
    return 0;
}

// MellingerControl._update_setpoint (:369-374), run every tick while no full-state command overrides the planner:
// TellState(self.state); UpdateTime(tick / 500); GetSetpoint(self.setpoint, self.state).  Also keeps state_t as the tick
// stored it (what _notifySetpointStop hands to TellState later).
template <typename T>
__device__ inline void hl_update_setpoint(HlState& h, const float pos[3], const float vel[3], float yaw_deg, int tick)
{
    for (int k = 0; k < 3; ++k) { h.st_pos[k] = pos[k]; h.st_vel[k] = vel[k]; }
    h.st_yaw_deg = yaw_deg;
    if (!h.active) return;
    Planner& p = h.plan;
    hl_tell_state(p, pos, vel, yaw_deg);
    p.t = (float)((double)tick / 500.0);
    const TrajEval ev = plan_current_goal<T>(p, p.t);
    if (p.state == PLAN_IDLE) hl_tell_state(p, pos, vel, yaw_deg);        // "on the ground": track the estimate
    if (!ev.valid) return;
    const float DEG = 180.0f / 3.14159265358979323846f;
    for (int k = 0; k < 3; ++k) {
        h.sp.pos[k] = ev.pos[k]; h.sp.vel[k] = ev.vel[k]; h.sp.acc[k] = ev.acc[k]; h.sp.rate_deg[k] = DEG * ev.omega[k];
        p.pos[k] = ev.pos[k]; p.vel[k] = ev.vel[k];                       // the commander stores its last setpoint
    }
    h.sp.yaw_deg = DEG * ev.yaw;
    h.sp.yaw_abs = 1;                                                     // mode.yaw = modeAbs, mode.quat = modeDisable
    p.yaw = ev.yaw;
}

// MellingerControl.reset (:137-150): crtpCommanderHighLevelInit(); TellState(state of the NOMINAL initial pose, E17).
template <typename T>
__device__ inline void hl_reset(const KParams<T>& P, const RaceParams<T>& RP, int i, HlState& h)
{
    h = HlState{};
    for (int k = 0; k < 3; ++k) h.st_pos[k] = (float)P.init_xyz[i][k];
    h.st_yaw_deg = (float)(RP.mel_prev_rpy[i][2] * (T)(180.0 / 3.14159265358979323846));
    hl_tell_state(h.plan, h.st_pos, h.st_vel, h.st_yaw_deg);
}

// One (Command, args) tuple (MultiRaceAviary.py:197-205 -> low_level_control, MellingerControl.py:32-57): the send* call
// queues the command and process_command_queue(args[-1]) (:292-303) pops it at once -- Stop(), UpdateTime(args[-1]) (the
// LAST argument, whatever it means for that command), then the command.  row: [adrp_command, args...] (16 floats).
// Returns true for ADRP_CMD_FULLSTATE: the caller then applies _sendFullStateCmd to row + 1.
template <typename T>
__device__ inline bool hl_command(HlState& h, const T* row)
{
    const int cmd = (int)row[0];
    const T* g = row + 1;
    if (cmd <= ADRP_CMD_NONE || cmd > ADRP_CMD_NOTIFY) return false;     // `else: continue`: no queue processing
    const int last_arg[11] = {0, 13, 1, 2, 2, 1, 2, 2, 0, 5, 0};
    Planner& p = h.plan;
    p.state = PLAN_IDLE;                                                  // crtpCommanderHighLevelStop()
    p.t = (float)g[last_arg[cmd]];                                        // crtpCommanderHighLevelUpdateTime(args[-1])
    switch (cmd) {
        case ADRP_CMD_FULLSTATE: h.active = 0; return true;               // full_state_cmd_override = True (:543)
        case ADRP_CMD_TAKEOFF: plan_takeoff_or_land(p, 0, (float)g[0], p.yaw, (float)g[1]); break;
        case ADRP_CMD_TAKEOFFYAW: plan_takeoff_or_land(p, 0, (float)g[0], (float)g[2], (float)g[1]); break;
        case ADRP_CMD_TAKEOFFVEL: hl_takeoff_vel(p, (float)g[0], (float)g[1], g[2] != T(0)); break;
        case ADRP_CMD_LAND: plan_takeoff_or_land(p, 1, (float)g[0], p.yaw, (float)g[1]); break;
        case ADRP_CMD_LANDYAW: plan_takeoff_or_land(p, 1, (float)g[0], (float)g[2], (float)g[1]); break;
        case ADRP_CMD_LANDVEL: hl_land_vel(p, (float)g[0], (float)g[1], g[2] != T(0)); break;
        case ADRP_CMD_STOP: p.state = PLAN_IDLE; break;
        case ADRP_CMD_GOTO: hl_go_to<T>(p, (float)g[0], (float)g[1], (float)g[2], (float)g[3], (float)g[4], g[5] != T(0)); break;
        case ADRP_CMD_NOTIFY: hl_tell_state(p, h.st_pos, h.st_vel, h.st_yaw_deg); break;   // _notifySetpointStop (:693-699)
    }
    h.active = 1;
    return false;
}

}  // namespace adrp
