// adrp_api.cu -- C ABI (include/adrp.h) over the fused step kernels.  Host side only: parameter
// derivation, buffer ownership, launches.  No arithmetic of the path runs on the host.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <mutex>
#include <new>
#include <string>

#include <vector>

#include "adrp_kernels.cuh"

using namespace adrp;

namespace {

// 16 independent FMA chains per thread, all operands in registers.
__global__ void fma_peak_kernel(float* sink, int iters, float a, float b)
{
    float x[16];
#pragma unroll
    for (int k = 0; k < 16; ++k) x[k] = (float)(threadIdx.x + k);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int k = 0; k < 16; ++k) x[k] = fmaf(x[k], a, b);
    }
    float s = 0;
#pragma unroll
    for (int k = 0; k < 16; ++k) s += x[k];
    if (s == 123.456f) sink[threadIdx.x & 1023] = s;
}

thread_local std::string g_create_error;
constexpr int kBlockHost = 128;          // = adrp::kBlock of the kernel translation units

struct HandleBase {
    adrp_config cfg;
    std::string err;
    int64_t launches = 0;
    int precision = 0;
    bool profiling = false;
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    double kernel_ms_sum = 0;
    long long kernel_ms_n = 0;
    bool ev_pending = false;
    // host staging for adrp_step_host
    void *h_action = nullptr, *h_obs = nullptr, *h_reward = nullptr;
    uint8_t *h_term = nullptr, *h_trunc = nullptr, *h_flags = nullptr;
    void *d_action = nullptr, *d_obs = nullptr, *d_reward = nullptr;
    uint8_t *d_term = nullptr, *d_trunc = nullptr, *d_flags = nullptr;
    cudaStream_t own_stream = nullptr, own_stream2 = nullptr;
    // Recorded on the caller's stream at the end of every device-path entry point (reset, set_state, get_state, step,
    // step_n); the host path's private streams wait on it before touching the handle's buffers, so that
    // set_state() / reset() followed by step_host() is ordered without a device-wide synchronisation.
    cudaEvent_t ev_order = nullptr;
    cudaEvent_t ev_join = nullptr;                  // step_host: the second private stream joins the first before the small copies
    bool order_pending = false;
    uint32_t rollout_step = 0;                      // ADRP_ACTIONS_PHILOX: control steps drawn so far (Philox counter word)
    virtual ~HandleBase() {}
};

template <typename T>
struct Handle : HandleBase {
    KParams<T> P;
    KBuffers<T> B;
    RaceParams<T> RP;
    RaceBuffers<T> RB;
    bool race = false;
};


// Every entry point runs on the handle's device, whatever the caller's current device is, and leaves the caller's
// device selected on return.
struct DeviceGuard {
    int prev = -1;
    explicit DeviceGuard(int device) { if (cudaGetDevice(&prev) == cudaSuccess && prev != device) cudaSetDevice(device); else prev = -1; }
    ~DeviceGuard() { if (prev >= 0) cudaSetDevice(prev); }
};

int next_pow2(int v) { int p = 1; while (p < v) p <<= 1; return p; }

// see HandleBase::ev_order
int mark_order(HandleBase* b, cudaStream_t s)
{
    if (!b->ev_order && cudaEventCreateWithFlags(&b->ev_order, cudaEventDisableTiming) != cudaSuccess) {
        b->err = "cudaEventCreateWithFlags failed"; return ADRP_ECUDA;
    }
    if (cudaEventRecord(b->ev_order, s) != cudaSuccess) { b->err = "cudaEventRecord failed"; return ADRP_ECUDA; }
    b->order_pending = true;
    return ADRP_OK;
}

int action_dim(int act)
{
    switch (act) {
    case ADRP_ACT_RPM_RAW: case ADRP_ACT_RPM: case ADRP_ACT_VEL: case ADRP_ACT_MEL: return 4;
    case ADRP_ACT_MEL_FULL: return 13;
    case ADRP_ACT_MEL_CMD: return 16;
    case ADRP_ACT_PID: return 3;
    case ADRP_ACT_ONE_D_RPM: case ADRP_ACT_ONE_D_PID: return 1;
    case ADRP_ACT_PID_FULL: return 12;
    }
    return 0;
}

// pybullet.getQuaternionFromEuler + the pose round trip, evaluated once on the host in double.
void host_init_quat(const double e[3], double q[4])
{
    double sp = sin(e[0] / 2), cp = cos(e[0] / 2), st = sin(e[1] / 2), ct = cos(e[1] / 2), ss = sin(e[2] / 2), cs = cos(e[2] / 2);
    double r[4] = {sp * ct * cs - cp * st * ss, cp * st * cs + sp * ct * ss, cp * ct * ss - sp * st * cs, cp * ct * cs + sp * st * ss};
    double n = sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2] + r[3] * r[3]);
    for (int k = 0; k < 4; ++k) r[k] /= n;
    double xx = r[0] * r[0], yy = r[1] * r[1], zz = r[2] * r[2], ww = r[3] * r[3];
    double pivot = 4 * ww > 1.0 ? r[3] : (xx < yy ? (yy < zz ? r[2] : r[1]) : (xx < zz ? r[2] : r[0]));
    double sgn = pivot < 0 ? -1.0 : 1.0;
    for (int k = 0; k < 4; ++k) q[k] = sgn * r[k];
}

template <typename T>
void derive_params(const adrp_config& c, KParams<T>& P)
{
    const adrp_drone_params& p = c.drone;
    memset(&P, 0, sizeof(P));
    P.N = c.num_envs; P.M = c.drones_per_env; P.Mpad = next_pow2(c.drones_per_env);
    P.S = c.pyb_freq / c.ctrl_freq; P.A = action_dim(c.action_type);
    P.B = (c.env_kind == ADRP_ENV_HOVER || c.env_kind == ADRP_ENV_MULTIHOVER) ? c.ctrl_freq / 2 : 0;   // BaseRLAviary.py:66
    P.O = c.env_kind == ADRP_ENV_CTRL ? (c.obs_columns == 12 ? 12 : 20) : 12 + P.B * P.A;
    if (c.env_kind == ADRP_ENV_RACE)      // MultiRaceAviary._observationSpace (:291-343)
        P.O = 12 + 5 * c.track.num_gates + 4 * c.track.num_obstacles + 1 + (c.race_mode == ADRP_RACE_COMPETE ? 6 * (c.drones_per_env - 1) : 0);
    P.physics = c.physics; P.drone_model = c.drone_model; P.action_type = c.action_type; P.env_kind = c.env_kind;
    P.race_mode = c.race_mode; P.auto_reset = c.auto_reset; P.pid_model = c.dslpid_model;
    P.Dpad = (long long)P.N * P.Mpad;
    P.env_begin = 0; P.env_end = P.N; P.obs_vec = 0; P.act_vec = 0;
    P.K = 1; P.act_step = P.obs_step = P.env_step = P.drone_step = P.noise_step = 0;
    P.act_rng = 0; P.rng_step0 = 0; P.rng_lo = (unsigned)(c.seed & 0xffffffffu); P.rng_hi = (unsigned)(c.seed >> 32);
    P.env_offset = c.env_offset;
    P.pyb_freq = c.pyb_freq;
    // step_counter / PYB_FREQ > EPISODE_LEN_SEC, evaluated like Python does (double division)
    int ts = 0;
    while ((double)ts / (double)c.pyb_freq <= c.episode_len_sec && ts < (1 << 30)) ts += P.S;
    P.trunc_steps = ts;
    // BaseAviary.__init__ derived constants (:116-128), in double, then rounded once to T
    const double GRAV = c.g * p.mass;
    const double HOVER = sqrt(GRAV / (4 * p.kf));
    const double MAXR = sqrt((p.thrust2weight * GRAV) / (4 * p.kf));
    const double MAXT = 4 * p.kf * MAXR * MAXR;
    P.mass = (T)p.mass; P.inv_mass = (T)(1.0 / p.mass); P.kf = (T)p.kf; P.km = (T)p.km;
    P.arm_x = (T)(c.drone_model == ADRP_CF2P ? p.arm : p.arm / sqrt(2.0));
    P.jx = (T)p.ixx; P.jy = (T)p.iyy; P.jz = (T)p.izz;
    P.jinvx = (T)(1.0 / p.ixx); P.jinvy = (T)(1.0 / p.iyy); P.jinvz = (T)(1.0 / p.izz);
    P.gravity = (T)GRAV; P.hover_rpm = (T)HOVER; P.max_rpm = (T)MAXR;
    P.gnd_eff_coeff = (T)p.gnd_eff_coeff; P.prop_radius = (T)p.prop_radius;
    P.gnd_eff_h_clip = (T)(0.25 * p.prop_radius * sqrt((15 * MAXR * MAXR * p.kf * p.gnd_eff_coeff) / MAXT));
    P.drag_xy = (T)p.drag_coeff_xy; P.drag_z = (T)p.drag_coeff_z;
    P.dw1 = (T)p.dw_coeff_1; P.dw2 = (T)p.dw_coeff_2; P.dw3 = (T)p.dw_coeff_3;
    {
        const double sc = 1.0 / sqrt(0.5 * 1.4426950408889634);
        const double c0 = p.dw_coeff_1 * (0.25 * p.prop_radius) * (0.25 * p.prop_radius);
        P.dwf_a = (T)(p.dw_coeff_2 * sc); P.dwf_b = (T)(p.dw_coeff_3 * sc);
        P.dwf_l2c = (T)(c0 > 0 ? log2(c0) : -1e30);
        P.half_dt = (T)(0.5 / c.pyb_freq); P.qr4 = (T)(0.25 * p.prop_radius);
    }
    {
        const double dt = 1.0 / c.pyb_freq, arm = c.drone_model == ADRP_CF2P ? p.arm : p.arm / sqrt(2.0);
        P.k_tx = (T)(dt / p.ixx * arm); P.k_ty = (T)(dt / p.iyy * arm);
        P.k_tz = (T)((c.drone_model == ADRP_RACE ? -1.0 : 1.0) * dt / p.izz);          // BaseAviary.py:850-852
        P.k_gx = (T)(dt / p.ixx * (p.izz - p.iyy)); P.k_gy = (T)(dt / p.iyy * (p.ixx - p.izz)); P.k_gz = (T)(dt / p.izz * (p.iyy - p.ixx));
        P.k_m = (T)(dt / p.mass); P.k_grav = (T)(dt * c.g);                            // GRAVITY / M = g
        P.k_drag_xy = (T)(dt / p.mass * p.drag_coeff_xy * (2.0 * M_PI / 60.0));
        P.k_drag_z = (T)(dt / p.mass * p.drag_coeff_z * (2.0 * M_PI / 60.0));
        P.k_lin = (T)(dt * p.lin_damping); P.k_ang = (T)(dt * p.ang_damping);
        P.plane_reach = (T)(sqrt(0.25 * p.collision_h * p.collision_h + p.collision_r * p.collision_r) * (1.0 + 1e-5) + 1e-6);
    }
    P.speed_limit = (T)(0.03 * p.max_speed_kmh * (1000.0 / 3600.0));
    for (int k = 0; k < 4; ++k) { P.prop_x[k] = (T)p.prop_x[k]; P.prop_y[k] = (T)p.prop_y[k]; }
    P.pyb_mass = (T)p.pyb_mass; P.pyb_jx = (T)p.pyb_ixx; P.pyb_jy = (T)p.pyb_iyy; P.pyb_jz = (T)p.pyb_izz;
    P.lin_damp = (T)p.lin_damping; P.ang_damp = (T)p.ang_damping;
    P.coll_r = (T)p.collision_r; P.coll_h = (T)p.collision_h;
    P.dt = (T)(1.0 / c.pyb_freq); P.ctrl_dt = (T)(1.0 / c.ctrl_freq); P.g = (T)c.g;
    P.pid_gravity = (T)(c.g * c.dslpid_drone.mass); P.pid_kf = (T)c.dslpid_drone.kf;
    P.hover_lim = (T)(c.env_kind == ADRP_ENV_HOVER ? 1.5 : 2.0);
    for (int i = 0; i < c.drones_per_env; ++i) {
        double q[4];
        host_init_quat(c.init_rpy[i], q);
        for (int k = 0; k < 4; ++k) P.init_quat[i][k] = (T)q[k];
        for (int k = 0; k < 3; ++k) {
            P.init_xyz[i][k] = (T)c.init_xyz[i][k]; P.init_vel[i][k] = (T)c.init_vel[i][k]; P.init_rpy[i][k] = (T)c.init_rpy[i][k];
            P.init_pqr[i][k] = (T)c.init_pqr[i][k]; P.target[i][k] = (T)c.target_pos[i][k];
        }
    }
}

#define CUDA_TRY(h, expr)                                                                  \
    do {                                                                                   \
        cudaError_t e_ = (expr);                                                           \
        if (e_ != cudaSuccess) {                                                           \
            (h)->err = std::string(#expr) + ": " + cudaGetErrorString(e_);                 \
            return ADRP_ECUDA;                                                             \
        }                                                                                  \
    } while (0)

// filter.c lpf2pSetCutoffFreq, evaluated once on the host in float like the firmware does.
void host_lpf2p(float sample_freq, float cutoff_freq, float out[5])
{
    const float fr = sample_freq / cutoff_freq;
    const float ohm = tanf(3.14159265358979323846f / fr);
    const float cc = 1.0f + 2.0f * cosf(3.14159265358979323846f / 4.0f) * ohm + ohm * ohm;
    out[0] = ohm * ohm / cc;
    out[1] = 2.0f * out[0];
    out[2] = out[0];
    out[3] = 2.0f * (ohm * ohm - 1.0f) / cc;
    out[4] = (1.0f - 2.0f * cosf(3.14159265358979323846f / 4.0f) * ohm + ohm * ohm) / cc;
}

template <typename T>
void derive_race(const adrp_config& c, const KParams<T>& P, RaceParams<T>& R)
{
    memset(&R, 0, sizeof(R));
    const adrp_track& t = c.track;
    const adrp_randomization& r = c.rnd;
    R.G = t.num_gates; R.NO = t.num_obstacles;
    R.TF = kGateFields * R.G + kObstFields * R.NO;
    R.R = 3 * R.G + 2 * R.NO + 10 * c.drones_per_env;
    R.compete = c.race_mode == ADRP_RACE_COMPETE;
    R.wrappers = c.wrappers;
    R.rnd_track = r.random_gates_obstacles; R.rnd_state = r.random_drone_state;
    R.rnd_inertia = r.random_drone_inertia; R.disturbances = r.disturbances;
    R.seed_lo = (unsigned)(c.seed & 0xffffffffu); R.seed_hi = (unsigned)(c.seed >> 32);
    R.env_offset = c.env_offset;
    for (int g = 0; g < R.G; ++g) {
        for (int k = 0; k < 4; ++k) R.gate_nom[g][k] = (T)t.gates[g][k];
        R.gate_type[g] = t.gate_type[g];
    }
    for (int o = 0; o < R.NO; ++o)
        for (int k = 0; k < 3; ++k) R.obst_nom[o][k] = (T)t.obstacles[o][k];
    for (int k = 0; k < 3; ++k) {
        R.bounds_hi[k] = (T)t.bounds_hi[k];
        R.pos_lo[k] = (T)r.pos_range[k][0]; R.pos_hi[k] = (T)r.pos_range[k][1];
        R.rot_lo[k] = (T)r.rot_range[k][0]; R.rot_hi[k] = (T)r.rot_range[k][1];
        R.wind_lo[k] = (T)r.wind_lo[k]; R.wind_hi[k] = (T)r.wind_hi[k];
    }
    R.contact_margin = (T)t.contact_margin; R.visibility = (T)t.visibility_range;
    R.gate_lo = (T)r.gate_range[0]; R.gate_hi = (T)r.gate_range[1];
    R.obst_lo = (T)r.obstacle_range[0]; R.obst_hi = (T)r.obstacle_range[1];
    const double* ranges[4] = {r.mass_range, r.ixx_range, r.iyy_range, r.izz_range};
    for (int k = 0; k < 4; ++k) { R.inertia_lo[k] = (T)ranges[k][0]; R.inertia_hi[k] = (T)ranges[k][1]; }
    R.noise_std = (T)r.action_noise_std;
    for (int i = 0; i < c.drones_per_env; ++i)
        for (int k = 0; k < 3; ++k) R.mel_prev_rpy[i][k] = (T)(c.init_rpy[i][k] * (M_PI / 180.0));   // MellingerControl.py:103-107
    R.mel_mass = (float)c.mel_mass;
    float lp[5];
    host_lpf2p(500.0f, 30.0f, lp);      // lpf2pInit(gyrolpf, FIRMWARE_FREQ, ACCEL_LPF_CUTOFF_FREQ): the swapped cutoffs of :116-117
    R.lpf_b0 = lp[0]; R.lpf_b1 = lp[1]; R.lpf_b2 = lp[2]; R.lpf_a1 = lp[3]; R.lpf_a2 = lp[4];
    // one firmware tick per physics substep while the episode lasts, plus the steps a caller may take after truncation
    const long long want = (long long)P.trunc_steps + 64LL * P.S + 1;          // beyond the table the kernel evaluates the float64 test itself
    R.lut_ticks = (int)(((want < (1LL << 22) ? want : (1LL << 22)) + 31) / 32) * 32;
}

template <typename T>
int alloc_buffers(Handle<T>* h)
{
    const KParams<T>& P = h->P;
    const size_t D = (size_t)P.Dpad;
    CUDA_TRY(h, cudaMalloc(&h->B.state, sizeof(T) * F_COUNT * D));
    CUDA_TRY(h, cudaMalloc(&h->B.pid, sizeof(T) * 9 * D));
    CUDA_TRY(h, cudaMalloc(&h->B.ring, sizeof(T) * (size_t)(P.B * P.A > 0 ? P.B * P.A : 1) * D));
    CUDA_TRY(h, cudaMalloc(&h->B.step_counter, sizeof(int32_t) * P.N));
    CUDA_TRY(h, cudaMalloc(&h->B.ring_head, sizeof(int32_t) * P.N));
    CUDA_TRY(h, cudaMemset(h->B.state, 0, sizeof(T) * F_COUNT * D));
    // BaseRLAviary.__init__ (:66-67) fills the action buffer with zeros ONCE; reset() keeps it
    CUDA_TRY(h, cudaMemset(h->B.ring, 0, sizeof(T) * (size_t)(P.B * P.A > 0 ? P.B * P.A : 1) * D));
    CUDA_TRY(h, cudaMemset(h->B.ring_head, 0, sizeof(int32_t) * P.N));
    if (h->race) {
        RaceBuffers<T>& rb = h->RB;
        const RaceParams<T>& R = h->RP;
        CUDA_TRY(h, cudaMalloc(&rb.track, sizeof(T) * (size_t)(R.TF > 0 ? R.TF : 1) * P.N));
        CUDA_TRY(h, cudaMalloc(&rb.body, sizeof(T) * 4 * D));
        CUDA_TRY(h, cudaMalloc(&rb.mel_t, sizeof(T) * MT_COUNT * D));
        CUDA_TRY(h, cudaMalloc(&rb.mel_f, sizeof(float) * MF_COUNT * D));
        CUDA_TRY(h, cudaMalloc(&rb.mel_i, sizeof(int32_t) * MI_COUNT * D));
        CUDA_TRY(h, cudaMalloc(&rb.gate_latch, sizeof(int32_t) * D));
        CUDA_TRY(h, cudaMalloc(&rb.episode, sizeof(int32_t) * P.N));
        CUDA_TRY(h, cudaMemset(rb.episode, 0, sizeof(int32_t) * P.N));
        CUDA_TRY(h, cudaMalloc(&rb.shape_gate, sizeof(int32_t) * P.N));
        CUDA_TRY(h, cudaMalloc(&rb.shape_vec, sizeof(T) * 6 * P.N));
        CUDA_TRY(h, cudaMemset(rb.shape_gate, 0xFF, sizeof(int32_t) * P.N));
        CUDA_TRY(h, cudaMemset(rb.shape_vec, 0, sizeof(T) * 6 * P.N));
        CUDA_TRY(h, cudaMemset(rb.mel_t, 0, sizeof(T) * MT_COUNT * D));
        CUDA_TRY(h, cudaMemset(rb.mel_f, 0, sizeof(float) * MF_COUNT * D));
        CUDA_TRY(h, cudaMemset(rb.mel_i, 0, sizeof(int32_t) * MI_COUNT * D));
        CUDA_TRY(h, cudaMemset(rb.gate_latch, 0, sizeof(int32_t) * D));
        rb.hl = nullptr;
        if (h->cfg.action_type == ADRP_ACT_MEL_CMD) {         // high-level commander state (planner, setpoint memory) per drone
            CUDA_TRY(h, cudaMalloc(&rb.hl, sizeof(uint32_t) * kHlWordsMax * D));
            CUDA_TRY(h, cudaMemset(rb.hl, 0, sizeof(uint32_t) * kHlWordsMax * D));
        }
        // sub-rate schedule (MellingerControl.py:395-411): bit t <=> (t/500 - (t-1)/500 > 0.002) in float64
        std::vector<uint32_t> lut((size_t)R.lut_ticks / 32, 0u);
        for (int t = 1; t < R.lut_ticks; ++t) {
            const volatile double a = (double)t / 500.0, b = (double)(t - 1) / 500.0;
            if (a - b > 0.002) lut[(size_t)t >> 5] |= 1u << (t & 31);
        }
        uint32_t* dl = nullptr;
        CUDA_TRY(h, cudaMalloc(&dl, sizeof(uint32_t) * lut.size()));
        CUDA_TRY(h, cudaMemcpy(dl, lut.data(), sizeof(uint32_t) * lut.size(), cudaMemcpyHostToDevice));
        rb.lut = dl;
        CUDA_TRY(h, Launch<T>::race_prepare(R, P));
    }
    return ADRP_OK;
}

template <typename T>
void free_buffers(Handle<T>* h)
{
    cudaFree(h->B.state); cudaFree(h->B.pid); cudaFree(h->B.ring); cudaFree(h->B.step_counter); cudaFree(h->B.ring_head);
    if (h->race) {
        RaceBuffers<T>& rb = h->RB;
        cudaFree(rb.track); cudaFree(rb.body); cudaFree(rb.mel_t); cudaFree(rb.mel_f); cudaFree(rb.mel_i);
        cudaFree(rb.hl);
        cudaFree(rb.gate_latch); cudaFree(rb.episode); cudaFree(rb.shape_gate); cudaFree(rb.shape_vec);
        cudaFree(const_cast<uint32_t*>(rb.lut));
    }
}

template <typename T>
int do_reset(Handle<T>* h, const uint8_t* mask, const T* draws, T* obs, cudaStream_t s)
{
    h->launches++;
    if (h->race) CUDA_TRY(h, Launch<T>::race_reset(h->P, h->RP, h->B, h->RB, mask, draws, obs, s));
    else CUDA_TRY(h, Launch<T>::reset(h->P, h->B, mask, obs, s));
    return ADRP_OK;
}

// K control steps in one launch (K = 1: adrp_step).  actions_per_step / outputs_per_step: the action (and noise) / output
// buffers hold one slice per control step, or one slice that every step reads / overwrites.
template <typename T>
int do_step(Handle<T>* h, const T* action, const T* noise, T* obs, T* reward, uint8_t* te, uint8_t* tr, uint8_t* fl, cudaStream_t s,
            int env_begin = 0, int env_end = -1, int K = 1, int actions_per_step = ADRP_ACTIONS_HELD, bool outputs_per_step = false)
{
    KParams<T> P = h->P;
    P.env_begin = env_begin; P.env_end = env_end < 0 ? P.N : env_end;
    P.K = K;
    const long long nd = (long long)P.N * P.M;
    if (actions_per_step == ADRP_ACTIONS_PER_STEP) P.act_step = nd * P.A;
    if (actions_per_step != ADRP_ACTIONS_HELD) P.noise_step = nd * P.S * 7;
    if (actions_per_step == ADRP_ACTIONS_PHILOX) { P.act_rng = 1; P.rng_step0 = h->rollout_step; h->rollout_step += (uint32_t)K; }
    if (outputs_per_step) { P.obs_step = nd * P.O; P.env_step = P.N; P.drone_step = nd; }
    P.obs_vec = sizeof(T) == 4 && (P.O % 4) == 0 && (P.A == 4 || P.env_kind == ADRP_ENV_CTRL) && ((uintptr_t)obs & 15) == 0;
    P.act_vec = sizeof(T) == 4 && P.A == 4 && ((uintptr_t)action & 15) == 0;
    if (h->profiling) {
        if (h->ev_pending) {          // harvest the previous launch's time before re-recording the events
            float ms = 0;
            CUDA_TRY(h, cudaEventSynchronize(h->ev1));
            CUDA_TRY(h, cudaEventElapsedTime(&ms, h->ev0, h->ev1));
            h->kernel_ms_sum += ms; h->kernel_ms_n++;
        }
        CUDA_TRY(h, cudaEventRecord(h->ev0, s));
    }
    h->launches++;
    if (h->race) CUDA_TRY(h, Launch<T>::race_step(P, h->RP, h->B, h->RB, action, noise, obs, reward, te, tr, fl, s));
    else CUDA_TRY(h, Launch<T>::step(P, h->B, action, obs, reward, te, tr, fl, s));
    if (h->profiling) { CUDA_TRY(h, cudaEventRecord(h->ev1, s)); h->ev_pending = true; }
    return ADRP_OK;
}

template <typename T>
int create_typed(const adrp_config* cfg, adrp_handle** out)
{
    Handle<T>* h = new (std::nothrow) Handle<T>();
    if (!h) { g_create_error = "out of host memory"; return ADRP_EINVAL; }
    h->cfg = *cfg;
    h->precision = cfg->precision;
    h->race = cfg->env_kind == ADRP_ENV_RACE;
    derive_params<T>(*cfg, h->P);
    if (h->race) derive_race<T>(*cfg, h->P, h->RP);
    int ndev = 0;
    cudaError_t e = cudaGetDeviceCount(&ndev);
    if (e != cudaSuccess || cfg->device < 0 || cfg->device >= ndev) {
        g_create_error = std::string("cudaSetDevice: ") + (e != cudaSuccess ? cudaGetErrorString(e) : "invalid device ordinal");
        delete h;
        return ADRP_ECUDA;
    }
    DeviceGuard dg(cfg->device);
    int rc = alloc_buffers(h);
    if (rc == ADRP_OK) rc = do_reset<T>(h, nullptr, nullptr, nullptr, 0);
    if (rc == ADRP_OK && cudaDeviceSynchronize() != cudaSuccess) { h->err = "reset kernel failed"; rc = ADRP_ECUDA; }
    if (rc != ADRP_OK) { g_create_error = h->err; free_buffers(h); delete h; return rc; }
    *out = reinterpret_cast<adrp_handle*>(static_cast<HandleBase*>(h));
    return ADRP_OK;
}

inline HandleBase* base(adrp_handle* h) { return reinterpret_cast<HandleBase*>(h); }
inline const HandleBase* base(const adrp_handle* h) { return reinterpret_cast<const HandleBase*>(h); }
template <typename T> Handle<T>* typed(adrp_handle* h) { return static_cast<Handle<T>*>(base(h)); }

int validate(const adrp_config* c)
{
    if (!c) { g_create_error = "config is NULL"; return ADRP_EINVAL; }
    if (c->abi_version != ADRP_ABI_VERSION) { g_create_error = "abi_version mismatch"; return ADRP_EABI; }
    if (c->num_envs < 1 || c->drones_per_env < 1 || c->drones_per_env > ADRP_MAX_DRONES) { g_create_error = "num_envs / drones_per_env out of range"; return ADRP_EINVAL; }
    if (c->ctrl_freq < 1 || c->pyb_freq % c->ctrl_freq != 0) { g_create_error = "[ERROR] in BaseAviary.__init__(), pyb_freq is not divisible by env_freq."; return ADRP_EINVAL; }
    if (c->precision != ADRP_F32 && c->precision != ADRP_F64) { g_create_error = "unknown precision"; return ADRP_EINVAL; }
    if (action_dim(c->action_type) == 0) { g_create_error = "unknown action_type"; return ADRP_EINVAL; }
    if ((c->env_kind == ADRP_ENV_RACE) != (c->action_type == ADRP_ACT_MEL || c->action_type == ADRP_ACT_MEL_FULL || c->action_type == ADRP_ACT_MEL_CMD)) { g_create_error = "ADRP_ACT_MEL / ADRP_ACT_MEL_FULL / ADRP_ACT_MEL_CMD (Mellinger setpoints and commands) are the race env's action types, and its only ones"; return ADRP_EINVAL; }
    if (c->wrappers & ~(ADRP_WRAP_OBS | ADRP_WRAP_REWARD)) { g_create_error = "unknown bit in wrappers"; return ADRP_EINVAL; }
    if (c->wrappers && c->env_kind != ADRP_ENV_RACE) { g_create_error = "wrappers (utils/wrapper.py) apply to the race env"; return ADRP_EINVAL; }
    if (c->env_kind == ADRP_ENV_RACE && (c->track.num_gates < 0 || c->track.num_gates > ADRP_MAX_GATES || c->track.num_obstacles < 0 || c->track.num_obstacles > ADRP_MAX_OBSTACLES)) { g_create_error = "too many gates / obstacles"; return ADRP_EINVAL; }
    if (c->obs_columns != 0 && !(c->env_kind == ADRP_ENV_CTRL && (c->obs_columns == 12 || c->obs_columns == 20))) { g_create_error = "obs_columns is 0, 12 or 20, and CtrlAviary-kind envs' only"; return ADRP_EINVAL; }
    if (c->env_kind == ADRP_ENV_HOVER && c->drones_per_env != 1) { g_create_error = "HoverAviary is single-drone"; return ADRP_EINVAL; }
    return ADRP_OK;
}

}  // namespace

extern "C" {

int adrp_abi_version(void) { return ADRP_ABI_VERSION; }
int adrp_config_size(void) { return (int)sizeof(adrp_config); }

int adrp_create(const adrp_config* cfg, adrp_handle** out)
{
    int rc = validate(cfg);
    if (rc != ADRP_OK) return rc;
    if (!out) { g_create_error = "out is NULL"; return ADRP_EINVAL; }
    return cfg->precision == ADRP_F64 ? create_typed<double>(cfg, out) : create_typed<float>(cfg, out);
}

int adrp_destroy(adrp_handle* h)
{
    if (!h) return ADRP_EINVAL;
    HandleBase* b = base(h);
    DeviceGuard dg(b->cfg.device);
    cudaDeviceSynchronize();
    if (b->precision == ADRP_F64) free_buffers(typed<double>(h)); else free_buffers(typed<float>(h));
    if (b->ev0) cudaEventDestroy(b->ev0);
    if (b->ev1) cudaEventDestroy(b->ev1);
    if (b->ev_order) cudaEventDestroy(b->ev_order);
    if (b->ev_join) cudaEventDestroy(b->ev_join);
    cudaFreeHost(b->h_action); cudaFreeHost(b->h_obs); cudaFreeHost(b->h_reward);
    cudaFreeHost(b->h_term); cudaFreeHost(b->h_trunc); cudaFreeHost(b->h_flags);
    cudaFree(b->d_action); cudaFree(b->d_obs); cudaFree(b->d_reward); cudaFree(b->d_term); cudaFree(b->d_trunc); cudaFree(b->d_flags);
    if (b->own_stream) cudaStreamDestroy(b->own_stream);
    if (b->own_stream2) cudaStreamDestroy(b->own_stream2);
    delete b;
    return ADRP_OK;
}

#define BY_PRECISION(h, expr64, expr32) (base(h)->precision == ADRP_F64 ? (expr64) : (expr32))

int adrp_obs_dim(const adrp_handle* h) { return !h ? 0 : BY_PRECISION(h, typed<double>(const_cast<adrp_handle*>(h))->P.O, typed<float>(const_cast<adrp_handle*>(h))->P.O); }
int adrp_action_dim(const adrp_handle* h) { return !h ? 0 : BY_PRECISION(h, typed<double>(const_cast<adrp_handle*>(h))->P.A, typed<float>(const_cast<adrp_handle*>(h))->P.A); }
int adrp_substeps(const adrp_handle* h) { return !h ? 0 : BY_PRECISION(h, typed<double>(const_cast<adrp_handle*>(h))->P.S, typed<float>(const_cast<adrp_handle*>(h))->P.S); }
int adrp_noise_dim(const adrp_handle*) { return 7; }
int adrp_reset_draws_dim(const adrp_handle* h)
{
    if (!h) return 0;
    const HandleBase* b = base(h);
    if (b->cfg.env_kind != ADRP_ENV_RACE) return 0;
    return 3 * b->cfg.track.num_gates + 2 * b->cfg.track.num_obstacles + 10 * b->cfg.drones_per_env;
}

int adrp_reset(adrp_handle* h, const uint8_t* env_mask, const void* draws, void* obs, void* stream)
{
    if (!h) return ADRP_EINVAL;
    DeviceGuard dg(base(h)->cfg.device);
    cudaStream_t s = (cudaStream_t)stream;
    const int rc = BY_PRECISION(h, do_reset<double>(typed<double>(h), env_mask, (const double*)draws, (double*)obs, s),
                                do_reset<float>(typed<float>(h), env_mask, (const float*)draws, (float*)obs, s));
    return rc != ADRP_OK ? rc : mark_order(base(h), s);
}

int adrp_set_state(adrp_handle* h, const void* state, void* stream)
{
    if (!h || !state) return ADRP_EINVAL;
    DeviceGuard dg(base(h)->cfg.device);
    cudaStream_t s = (cudaStream_t)stream;
    base(h)->launches++;
    if (base(h)->precision == ADRP_F64) {
        auto* t = typed<double>(h);
        CUDA_TRY(t, Launch<double>::set_state(t->P, t->B, (const double*)state, s));
    } else {
        auto* t = typed<float>(h);
        CUDA_TRY(t, Launch<float>::set_state(t->P, t->B, (const float*)state, s));
    }
    return mark_order(base(h), s);
}

int adrp_get_state(adrp_handle* h, void* state, void* stream)
{
    if (!h || !state) return ADRP_EINVAL;
    DeviceGuard dg(base(h)->cfg.device);
    cudaStream_t s = (cudaStream_t)stream;
    base(h)->launches++;
    if (base(h)->precision == ADRP_F64) {
        auto* t = typed<double>(h);
        CUDA_TRY(t, Launch<double>::get_state(t->P, t->B, (double*)state, s));
    } else {
        auto* t = typed<float>(h);
        CUDA_TRY(t, Launch<float>::get_state(t->P, t->B, (float*)state, s));
    }
    return mark_order(base(h), s);
}

int adrp_step(adrp_handle* h, const void* action, const void* noise, void* obs, void* reward,
              uint8_t* terminated, uint8_t* truncated, uint8_t* drone_flags, void* stream)
{
    if (!h || !action || !obs || !reward || !terminated || !truncated) {
        if (h) base(h)->err = "adrp_step: NULL buffer";
        return ADRP_EINVAL;
    }
    DeviceGuard dg(base(h)->cfg.device);
    cudaStream_t s = (cudaStream_t)stream;
    const int rc = BY_PRECISION(h, do_step<double>(typed<double>(h), (const double*)action, (const double*)noise, (double*)obs, (double*)reward, terminated, truncated, drone_flags, s),
                                do_step<float>(typed<float>(h), (const float*)action, (const float*)noise, (float*)obs, (float*)reward, terminated, truncated, drone_flags, s));
    return rc != ADRP_OK ? rc : mark_order(base(h), s);
}

int adrp_set_rollout_step(adrp_handle* h, uint32_t step)
{
    if (!h) return ADRP_EINVAL;
    base(h)->rollout_step = step;
    return ADRP_OK;
}

int adrp_step_n(adrp_handle* h, int n_steps, int actions_per_step, int outputs_per_step, const void* action, const void* noise,
                void* obs, void* reward, uint8_t* terminated, uint8_t* truncated, uint8_t* drone_flags, void* stream)
{
    const bool drawn = actions_per_step == ADRP_ACTIONS_PHILOX;
    if (!h || (!action && !drawn) || !obs || !reward || !terminated || !truncated || n_steps < 1) {
        if (h) base(h)->err = "adrp_step_n: NULL buffer or n_steps < 1";
        return ADRP_EINVAL;
    }
    if (actions_per_step < ADRP_ACTIONS_HELD || actions_per_step > ADRP_ACTIONS_PHILOX) {
        base(h)->err = "adrp_step_n: actions_per_step is ADRP_ACTIONS_HELD, ADRP_ACTIONS_PER_STEP or ADRP_ACTIONS_PHILOX";
        return ADRP_EINVAL;
    }
    if (drawn && action_dim(base(h)->cfg.action_type) > 4) {
        base(h)->err = "adrp_step_n: ADRP_ACTIONS_PHILOX draws the Box(-1, 1) actions of at most 4 components (not ADRP_ACT_PID_FULL / ADRP_ACT_MEL_FULL / ADRP_ACT_MEL_CMD rows)";
        return ADRP_EUNSUPPORTED;
    }
    DeviceGuard dg(base(h)->cfg.device);
    cudaStream_t s = (cudaStream_t)stream;
    const int aps = actions_per_step;
    const bool ops = outputs_per_step != 0;
    const int rc = BY_PRECISION(h, do_step<double>(typed<double>(h), (const double*)action, (const double*)noise, (double*)obs, (double*)reward, terminated, truncated, drone_flags, s, 0, -1, n_steps, aps, ops),
                                do_step<float>(typed<float>(h), (const float*)action, (const float*)noise, (float*)obs, (float*)reward, terminated, truncated, drone_flags, s, 0, -1, n_steps, aps, ops));
    return rc != ADRP_OK ? rc : mark_order(base(h), s);
}

}  // extern "C"

namespace {

template <typename T>
int ensure_staging(Handle<T>* h)
{
    if (h->own_stream) return ADRP_OK;
    const KParams<T>& P = h->P;
    const size_t nd = (size_t)P.N * P.M;
    CUDA_TRY(h, cudaStreamCreateWithFlags(&h->own_stream, cudaStreamNonBlocking));
    CUDA_TRY(h, cudaStreamCreateWithFlags(&h->own_stream2, cudaStreamNonBlocking));
    CUDA_TRY(h, cudaMallocHost(&h->h_action, sizeof(T) * nd * P.A));
    CUDA_TRY(h, cudaMallocHost(&h->h_obs, sizeof(T) * nd * P.O));
    CUDA_TRY(h, cudaMallocHost(&h->h_reward, sizeof(T) * P.N));
    CUDA_TRY(h, cudaMallocHost(&h->h_term, P.N));
    CUDA_TRY(h, cudaMallocHost(&h->h_trunc, P.N));
    CUDA_TRY(h, cudaMallocHost(&h->h_flags, nd));
    CUDA_TRY(h, cudaMalloc(&h->d_action, sizeof(T) * nd * P.A));
    CUDA_TRY(h, cudaMalloc(&h->d_obs, sizeof(T) * nd * P.O));
    CUDA_TRY(h, cudaMalloc(&h->d_reward, sizeof(T) * P.N));
    CUDA_TRY(h, cudaMalloc(&h->d_term, P.N));
    CUDA_TRY(h, cudaMalloc(&h->d_trunc, P.N));
    CUDA_TRY(h, cudaMalloc(&h->d_flags, nd));
    return ADRP_OK;
}

bool is_pinned(const void* p)
{
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, p) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeHost;
}

// Host-buffer step (the call a NumPy user of the reference makes).  The env range is cut into chunks that
// alternate between two streams: chunk c+1's action upload and kernel overlap chunk c's observation download
// (envs are independent, so a chunk is a complete step of its envs).  Page-locked caller buffers (cudaHostAlloc /
// cudaHostRegister, e.g. torch pin_memory) are DMA'd directly; pageable ones go through the handle's pinned staging.
template <typename T>
int do_step_host(Handle<T>* h, const void* action, void* obs, void* reward, uint8_t* te, uint8_t* tr, uint8_t* fl)
{
    int rc = ensure_staging(h);
    if (rc != ADRP_OK) return rc;
    const KParams<T>& P = h->P;
    const size_t nd = (size_t)P.N * P.M;
    const bool pin_a = is_pinned(action), pin_o = is_pinned(obs);
    const bool pin_r = is_pinned(reward), pin_te = is_pinned(te), pin_tr = is_pinned(tr), pin_fl = fl && is_pinned(fl);
    const T* a_src = pin_a ? (const T*)action : (const T*)h->h_action;
    T* o_dst = pin_o ? (T*)obs : (T*)h->h_obs;
    if (!pin_a) memcpy(h->h_action, action, sizeof(T) * nd * P.A);
    if (h->order_pending) {                                  // earlier device-path work on this handle (reset, set_state, step)
        CUDA_TRY(h, cudaStreamWaitEvent(h->own_stream, h->ev_order, 0));
        CUDA_TRY(h, cudaStreamWaitEvent(h->own_stream2, h->ev_order, 0));
        h->order_pending = false;
    }
    if (!h->ev_join) CUDA_TRY(h, cudaEventCreateWithFlags(&h->ev_join, cudaEventDisableTiming));
    const int epb = kBlockHost / P.Mpad;                                   // envs per thread block
    int chunks = (int)(nd * P.O * sizeof(T) / (8u << 20));                  // ~8 MiB of observations per chunk
    chunks = chunks < 1 ? 1 : (chunks > 8 ? 8 : chunks);
    int per = ((P.N + chunks - 1) / chunks + epb - 1) / epb * epb;
    bool used2 = false;
    for (int c = 0, e0 = 0; e0 < P.N; ++c, e0 += per) {
        const int e1 = e0 + per < P.N ? e0 + per : P.N;
        cudaStream_t s = c & 1 ? h->own_stream2 : h->own_stream;
        used2 |= (c & 1) != 0;
        const size_t r0 = (size_t)e0 * P.M, rn = (size_t)(e1 - e0) * P.M;
        CUDA_TRY(h, cudaMemcpyAsync((T*)h->d_action + r0 * P.A, a_src + r0 * P.A, sizeof(T) * rn * P.A, cudaMemcpyHostToDevice, s));
        rc = do_step<T>(h, (const T*)h->d_action, (const T*)nullptr, (T*)h->d_obs, (T*)h->d_reward, h->d_term, h->d_trunc, h->d_flags, s, e0, e1);
        if (rc != ADRP_OK) return rc;
        CUDA_TRY(h, cudaMemcpyAsync(o_dst + r0 * P.O, (T*)h->d_obs + r0 * P.O, sizeof(T) * rn * P.O, cudaMemcpyDeviceToHost, s));
    }
    // the small per-env / per-drone outputs: ONE copy each for the whole batch, behind the kernels of both streams (per
    // chunk they were 4 extra cudaMemcpyAsync calls each: 32 launches of the copy engine per step for 2 % of the bytes)
    if (used2) {
        CUDA_TRY(h, cudaEventRecord(h->ev_join, h->own_stream2));
        CUDA_TRY(h, cudaStreamWaitEvent(h->own_stream, h->ev_join, 0));
    }
    cudaStream_t s1 = h->own_stream;
    CUDA_TRY(h, cudaMemcpyAsync(pin_r ? reward : h->h_reward, h->d_reward, sizeof(T) * P.N, cudaMemcpyDeviceToHost, s1));
    CUDA_TRY(h, cudaMemcpyAsync(pin_te ? te : h->h_term, h->d_term, P.N, cudaMemcpyDeviceToHost, s1));
    CUDA_TRY(h, cudaMemcpyAsync(pin_tr ? tr : h->h_trunc, h->d_trunc, P.N, cudaMemcpyDeviceToHost, s1));
    if (fl) CUDA_TRY(h, cudaMemcpyAsync(pin_fl ? fl : h->h_flags, h->d_flags, nd, cudaMemcpyDeviceToHost, s1));
    CUDA_TRY(h, cudaStreamSynchronize(h->own_stream2));
    CUDA_TRY(h, cudaStreamSynchronize(h->own_stream));
    if (!pin_o) memcpy(obs, h->h_obs, sizeof(T) * nd * P.O);
    if (!pin_r) memcpy(reward, h->h_reward, sizeof(T) * P.N);
    if (!pin_te) memcpy(te, h->h_term, P.N);
    if (!pin_tr) memcpy(tr, h->h_trunc, P.N);
    if (fl && !pin_fl) memcpy(fl, h->h_flags, nd);
    return ADRP_OK;
}

}  // namespace

extern "C" {

int adrp_step_host(adrp_handle* h, const void* action, void* obs, void* reward, uint8_t* terminated,
                   uint8_t* truncated, uint8_t* drone_flags)
{
    if (!h || !action || !obs || !reward || !terminated || !truncated) {
        if (h) base(h)->err = "adrp_step_host: NULL buffer";
        return ADRP_EINVAL;
    }
    DeviceGuard dg(base(h)->cfg.device);
    return BY_PRECISION(h, do_step_host<double>(typed<double>(h), action, obs, reward, terminated, truncated, drone_flags),
                        do_step_host<float>(typed<float>(h), action, obs, reward, terminated, truncated, drone_flags));
}

int adrp_dslpid_control(int precision, int drone_model, const adrp_drone_params* drone, double g, int n,
                        double control_timestep, const void* pos, const void* quat, const void* vel,
                        const void* target_pos, const void* target_rpy, const void* target_vel,
                        const void* target_rates, void* ctrl_state, void* rpm, void* pos_e, void* yaw_e, void* stream)
{
    if (!drone || n < 0 || !pos || !quat || !vel || !target_pos || !ctrl_state || !rpm) return ADRP_EINVAL;
    if (drone_model != ADRP_CF2X && drone_model != ADRP_CF2P) return ADRP_EUNSUPPORTED;   // DSLPIDControl.py:34-36
    if (n == 0) return ADRP_OK;
    // run on the device that owns the buffers, whatever the caller's current device is
    cudaPointerAttributes at;
    if (cudaPointerGetAttributes(&at, rpm) != cudaSuccess || at.type != cudaMemoryTypeDevice) { cudaGetLastError(); return ADRP_EINVAL; }
    DeviceGuard dg(at.device);
    cudaStream_t s = (cudaStream_t)stream;
    cudaError_t e;
    if (precision == ADRP_F64)
        e = Launch<double>::dslpid_rows(g * drone->mass, drone->kf, drone_model, n, control_timestep,
            (const double*)pos, (const double*)quat, (const double*)vel, (const double*)target_pos, (const double*)target_rpy,
            (const double*)target_vel, (const double*)target_rates, (double*)ctrl_state, (double*)rpm, (double*)pos_e, (double*)yaw_e, s);
    else
        e = Launch<float>::dslpid_rows((float)(g * drone->mass), (float)drone->kf, drone_model, n, (float)control_timestep,
            (const float*)pos, (const float*)quat, (const float*)vel, (const float*)target_pos, (const float*)target_rpy,
            (const float*)target_vel, (const float*)target_rates, (float*)ctrl_state, (float*)rpm, (float*)pos_e, (float*)yaw_e, s);
    return e == cudaSuccess ? ADRP_OK : ADRP_ECUDA;
}

int64_t adrp_launch_count(const adrp_handle* h) { return h ? base(h)->launches : 0; }

int adrp_set_profiling(adrp_handle* h, int enabled)
{
    if (!h) return ADRP_EINVAL;
    HandleBase* b = base(h);
    DeviceGuard dg(b->cfg.device);
    if (enabled && !b->ev0) {
        CUDA_TRY(b, cudaEventCreate(&b->ev0));
        CUDA_TRY(b, cudaEventCreate(&b->ev1));
    }
    b->profiling = enabled != 0;
    b->ev_pending = false;
    b->kernel_ms_sum = 0; b->kernel_ms_n = 0;
    return ADRP_OK;
}

double adrp_step_kernel_ms(adrp_handle* h, int reset)
{
    if (!h) return -1.0;
    HandleBase* b = base(h);
    DeviceGuard dg(b->cfg.device);
    if (b->ev_pending) {
        float ms = 0;
        if (cudaEventSynchronize(b->ev1) == cudaSuccess && cudaEventElapsedTime(&ms, b->ev0, b->ev1) == cudaSuccess) {
            b->kernel_ms_sum += ms; b->kernel_ms_n++;
        }
        b->ev_pending = false;
    }
    const double avg = b->kernel_ms_n ? b->kernel_ms_sum / (double)b->kernel_ms_n : -1.0;
    if (reset) { b->kernel_ms_sum = 0; b->kernel_ms_n = 0; }
    return avg;
}

double adrp_fp32_peak_tflops(int device)
{
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || device < 0 || device >= ndev) return -1.0;
    DeviceGuard dg(device);                               // (the caller's device is selected again on return)
    int sms = 0;
    if (cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, device) != cudaSuccess) return -1.0;
    float* sink = nullptr;
    if (cudaMalloc(&sink, sizeof(float) * 1024) != cudaSuccess) return -1.0;
    const int iters = 4096, blocks = sms * 8, threads = 256;
    cudaEvent_t e0 = nullptr, e1 = nullptr;
    if (cudaEventCreate(&e0) != cudaSuccess || cudaEventCreate(&e1) != cudaSuccess) {
        if (e0) cudaEventDestroy(e0);
        cudaFree(sink);
        return -1.0;
    }
    double best = -1.0;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        fma_peak_kernel<<<blocks, threads>>>(sink, iters, 1.0001f, 0.9999f);
        cudaEventRecord(e1);
        if (cudaEventSynchronize(e1) != cudaSuccess) { best = -1.0; break; }
        float ms = 0;
        cudaEventElapsedTime(&ms, e0, e1);
        const double flops = 2.0 * 16.0 * (double)iters * (double)blocks * threads;
        const double tf = flops / (ms * 1e-3) / 1e12;
        if (rep > 0 && tf > best) best = tf;
    }
    cudaEventDestroy(e0); cudaEventDestroy(e1); cudaFree(sink);
    return best;
}

const char* adrp_last_error(const adrp_handle* h) { return h ? base(h)->err.c_str() : g_create_error.c_str(); }

}  // extern "C"

// ---------------------------------------------------------------------------------------------------------------------
// adrp_step_rows_static (include/adrp.h): a control step without a handle and without any allocation.  No counterpart in
// the reference.
//
// The SoA working state of step_kernel lives in this file-scope device array.  In HOST code the name `g_static_scratch`
// is only a stand-in for the symbol, NOT a device address (a kernel handed it would fault, or silently use host memory
// where the GPU can reach it): the address comes from cudaGetSymbolAddress, once per device, with that device current.
static __device__ __align__(256) unsigned char g_static_scratch[ADRP_STATIC_SCRATCH_BYTES];

namespace {

constexpr int kMaxStaticDevices = 64;
std::mutex g_static_mu;
unsigned char* g_static_ptr[kMaxStaticDevices] = {};      // device address of g_static_scratch on each device

// Must be called with `device` current (DeviceGuard).
int static_scratch_ptr(int device, unsigned char** out)
{
    if (device < 0 || device >= kMaxStaticDevices) { g_create_error = "adrp_step_rows_static: device ordinal out of range"; return ADRP_EINVAL; }
    std::lock_guard<std::mutex> lock(g_static_mu);
    if (!g_static_ptr[device]) {
        void* p = nullptr;
        const cudaError_t e = cudaGetSymbolAddress(&p, g_static_scratch);
        if (e != cudaSuccess || !p) {
            g_create_error = std::string("adrp_step_rows_static: cudaGetSymbolAddress: ") + cudaGetErrorString(e);
            return ADRP_ECUDA;
        }
        g_static_ptr[device] = (unsigned char*)p;
    }
    *out = g_static_ptr[device];
    return ADRP_OK;
}

template <typename T>
int step_rows_static(const adrp_config* cfg, const T* state_in, const T* action, T* state_out, T* obs, T* reward,
                     uint8_t* te, uint8_t* tr, cudaStream_t s)
{
    KParams<T> P;
    derive_params<T>(*cfg, P);
    const size_t state_bytes = sizeof(T) * (size_t)F_COUNT * (size_t)P.Dpad;
    const size_t counter_off = (state_bytes + 255) & ~(size_t)255;
    const size_t need = counter_off + sizeof(int32_t) * (size_t)P.N;
    if (need > (size_t)ADRP_STATIC_SCRATCH_BYTES) {
        char msg[256];
        snprintf(msg, sizeof(msg), "adrp_step_rows_static: %d envs x %d lanes need %zu bytes of scratch, the static array holds %u "
                 "(ADRP_STATIC_SCRATCH_BYTES): use a handle (adrp_create) for batches of this size", P.N, P.Mpad, need,
                 (unsigned)ADRP_STATIC_SCRATCH_BYTES);
        g_create_error = msg;
        return ADRP_EINVAL;
    }
    unsigned char* base_ptr = nullptr;
    const int rc = static_scratch_ptr(cfg->device, &base_ptr);
    if (rc != ADRP_OK) return rc;
    KBuffers<T> B;
    B.state = reinterpret_cast<T*>(base_ptr);
    B.pid = nullptr; B.ring = nullptr; B.ring_head = nullptr;        // never touched by RPM-type actions on a CtrlAviary-kind env
    B.step_counter = reinterpret_cast<int32_t*>(base_ptr + counter_off);
    auto fail = [](const char* what, cudaError_t e) { g_create_error = std::string("adrp_step_rows_static: ") + what + ": " + cudaGetErrorString(e); return ADRP_ECUDA; };
    cudaError_t e;
    // the array keeps the last call's contents: the step counters must start at zero, the state part is overwritten below
    if ((e = cudaMemsetAsync(B.step_counter, 0, sizeof(int32_t) * (size_t)P.N, s)) != cudaSuccess) return fail("cudaMemsetAsync", e);
    if ((e = Launch<T>::set_state(P, B, state_in, s)) != cudaSuccess) return fail("set_state", e);
    P.obs_vec = sizeof(T) == 4 && (P.O % 4) == 0 && ((uintptr_t)obs & 15) == 0;          // as do_step decides it
    P.act_vec = sizeof(T) == 4 && P.A == 4 && ((uintptr_t)action & 15) == 0;
    if ((e = Launch<T>::step(P, B, action, obs, reward, te, tr, nullptr, s)) != cudaSuccess) return fail("step", e);
    P.obs_vec = P.act_vec = 0;
    if ((e = Launch<T>::get_state(P, B, state_out, s)) != cudaSuccess) return fail("get_state", e);
    return ADRP_OK;
}

}  // namespace

extern "C" int adrp_step_rows_static(const adrp_config* cfg, const void* state_in, const void* action, void* state_out,
                                     void* obs, void* reward, uint8_t* terminated, uint8_t* truncated, void* stream)
{
    int rc = validate(cfg);
    if (rc != ADRP_OK) return rc;
    if (!state_in || !action || !state_out || !obs || !reward || !terminated || !truncated) { g_create_error = "adrp_step_rows_static: NULL buffer"; return ADRP_EINVAL; }
    if (cfg->env_kind != ADRP_ENV_CTRL || cfg->action_type != ADRP_ACT_RPM_RAW) {
        g_create_error = "adrp_step_rows_static: env_kind must be ADRP_ENV_CTRL and action_type ADRP_ACT_RPM_RAW";
        return ADRP_EUNSUPPORTED;
    }
    int ndev = 0;
    if (cudaGetDeviceCount(&ndev) != cudaSuccess || cfg->device < 0 || cfg->device >= ndev) { g_create_error = "adrp_step_rows_static: invalid device ordinal"; return ADRP_ECUDA; }
    DeviceGuard dg(cfg->device);
    cudaStream_t s = (cudaStream_t)stream;
    return cfg->precision == ADRP_F64
        ? step_rows_static<double>(cfg, (const double*)state_in, (const double*)action, (double*)state_out, (double*)obs, (double*)reward, terminated, truncated, s)
        : step_rows_static<float>(cfg, (const float*)state_in, (const float*)action, (float*)state_out, (float*)obs, (float*)reward, terminated, truncated, s);
}
