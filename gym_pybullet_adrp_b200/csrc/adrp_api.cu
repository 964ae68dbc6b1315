// adrp_api.cu -- C ABI (include/adrp.h) over the fused step kernels.  Host side only: parameter
// derivation, buffer ownership, launches.  No arithmetic of the path runs on the host.
#include <cuda_runtime.h>

#include <cmath>
#include <cstdio>
#include <cstring>
#include <new>
#include <string>

#include "adrp_step.cuh"

using namespace adrp;

namespace {

thread_local std::string g_create_error;

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
    cudaStream_t own_stream = nullptr;
    virtual ~HandleBase() {}
};

template <typename T>
struct Handle : HandleBase {
    KParams<T> P;
    KBuffers<T> B;
};

int next_pow2(int v) { int p = 1; while (p < v) p <<= 1; return p; }

int action_dim(int act)
{
    switch (act) {
    case ADRP_ACT_RPM_RAW: case ADRP_ACT_RPM: case ADRP_ACT_VEL: case ADRP_ACT_MEL: return 4;
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
    P.O = c.env_kind == ADRP_ENV_CTRL ? 20 : 12 + P.B * P.A;
    P.physics = c.physics; P.drone_model = c.drone_model; P.action_type = c.action_type; P.env_kind = c.env_kind;
    P.race_mode = c.race_mode; P.auto_reset = c.auto_reset; P.pid_model = c.dslpid_model;
    P.Dpad = (long long)P.N * P.Mpad;
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
    P.mass = (T)p.mass; P.kf = (T)p.kf; P.km = (T)p.km;
    P.arm_x = (T)(c.drone_model == ADRP_CF2P ? p.arm : p.arm / sqrt(2.0));
    P.jx = (T)p.ixx; P.jy = (T)p.iyy; P.jz = (T)p.izz;
    P.jinvx = (T)(1.0 / p.ixx); P.jinvy = (T)(1.0 / p.iyy); P.jinvz = (T)(1.0 / p.izz);
    P.gravity = (T)GRAV; P.hover_rpm = (T)HOVER; P.max_rpm = (T)MAXR;
    P.gnd_eff_coeff = (T)p.gnd_eff_coeff; P.prop_radius = (T)p.prop_radius;
    P.gnd_eff_h_clip = (T)(0.25 * p.prop_radius * sqrt((15 * MAXR * MAXR * p.kf * p.gnd_eff_coeff) / MAXT));
    P.drag_xy = (T)p.drag_coeff_xy; P.drag_z = (T)p.drag_coeff_z;
    P.dw1 = (T)p.dw_coeff_1; P.dw2 = (T)p.dw_coeff_2; P.dw3 = (T)p.dw_coeff_3;
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
            P.init_xyz[i][k] = (T)c.init_xyz[i][k]; P.init_vel[i][k] = (T)c.init_vel[i][k];
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
    return ADRP_OK;
}

template <typename T>
void free_buffers(Handle<T>* h)
{
    cudaFree(h->B.state); cudaFree(h->B.pid); cudaFree(h->B.ring); cudaFree(h->B.step_counter); cudaFree(h->B.ring_head);
}

inline unsigned grid_for(long long D) { return (unsigned)((D + kBlock - 1) / kBlock); }

template <typename T>
int do_reset(Handle<T>* h, const uint8_t* mask, T* obs, cudaStream_t s)
{
    reset_kernel<T><<<grid_for(h->P.Dpad), kBlock, 0, s>>>(h->P, h->B, mask, obs);
    h->launches++;
    CUDA_TRY(h, cudaGetLastError());
    return ADRP_OK;
}

template <typename T>
int do_step(Handle<T>* h, const T* action, T* obs, T* reward, uint8_t* te, uint8_t* tr, uint8_t* fl, cudaStream_t s)
{
    const KParams<T>& P = h->P;
    const unsigned grid = grid_for(P.Dpad);
    const bool pid = P.action_type == ADRP_ACT_PID || P.action_type == ADRP_ACT_ONE_D_PID ||
                     P.action_type == ADRP_ACT_VEL || P.action_type == ADRP_ACT_PID_FULL;
    if (h->profiling) {
        if (h->ev_pending) {          // harvest the previous launch's time before re-recording the events
            float ms = 0;
            CUDA_TRY(h, cudaEventSynchronize(h->ev1));
            CUDA_TRY(h, cudaEventElapsedTime(&ms, h->ev0, h->ev1));
            h->kernel_ms_sum += ms; h->kernel_ms_n++;
        }
        CUDA_TRY(h, cudaEventRecord(h->ev0, s));
    }
    if (pid) step_kernel<T, -1, 1><<<grid, kBlock, 0, s>>>(P, h->B, action, obs, reward, te, tr, fl);
    else step_kernel<T, -1, 0><<<grid, kBlock, 0, s>>>(P, h->B, action, obs, reward, te, tr, fl);
    h->launches++;
    if (h->profiling) { CUDA_TRY(h, cudaEventRecord(h->ev1, s)); h->ev_pending = true; }
    CUDA_TRY(h, cudaGetLastError());
    return ADRP_OK;
}

template <typename T>
int create_typed(const adrp_config* cfg, adrp_handle** out)
{
    Handle<T>* h = new (std::nothrow) Handle<T>();
    if (!h) { g_create_error = "out of host memory"; return ADRP_EINVAL; }
    h->cfg = *cfg;
    h->precision = cfg->precision;
    derive_params<T>(*cfg, h->P);
    cudaError_t e = cudaSetDevice(cfg->device);
    if (e != cudaSuccess) { g_create_error = std::string("cudaSetDevice: ") + cudaGetErrorString(e); delete h; return ADRP_ECUDA; }
    int rc = alloc_buffers(h);
    if (rc == ADRP_OK) rc = do_reset<T>(h, nullptr, nullptr, 0);
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
    if (c->env_kind == ADRP_ENV_RACE || c->action_type == ADRP_ACT_MEL) { g_create_error = "race env / Mellinger: not in this build"; return ADRP_EUNSUPPORTED; }
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
    cudaDeviceSynchronize();
    if (b->precision == ADRP_F64) free_buffers(typed<double>(h)); else free_buffers(typed<float>(h));
    if (b->ev0) cudaEventDestroy(b->ev0);
    if (b->ev1) cudaEventDestroy(b->ev1);
    cudaFreeHost(b->h_action); cudaFreeHost(b->h_obs); cudaFreeHost(b->h_reward);
    cudaFreeHost(b->h_term); cudaFreeHost(b->h_trunc); cudaFreeHost(b->h_flags);
    cudaFree(b->d_action); cudaFree(b->d_obs); cudaFree(b->d_reward); cudaFree(b->d_term); cudaFree(b->d_trunc); cudaFree(b->d_flags);
    if (b->own_stream) cudaStreamDestroy(b->own_stream);
    delete b;
    return ADRP_OK;
}

#define BY_PRECISION(h, expr64, expr32) (base(h)->precision == ADRP_F64 ? (expr64) : (expr32))

int adrp_obs_dim(const adrp_handle* h) { return BY_PRECISION(h, typed<double>(const_cast<adrp_handle*>(h))->P.O, typed<float>(const_cast<adrp_handle*>(h))->P.O); }
int adrp_action_dim(const adrp_handle* h) { return BY_PRECISION(h, typed<double>(const_cast<adrp_handle*>(h))->P.A, typed<float>(const_cast<adrp_handle*>(h))->P.A); }
int adrp_substeps(const adrp_handle* h) { return BY_PRECISION(h, typed<double>(const_cast<adrp_handle*>(h))->P.S, typed<float>(const_cast<adrp_handle*>(h))->P.S); }
int adrp_noise_dim(const adrp_handle*) { return 7; }
int adrp_reset_draws_dim(const adrp_handle*) { return 0; }

int adrp_reset(adrp_handle* h, const uint8_t* env_mask, const void* draws, void* obs, void* stream)
{
    (void)draws;
    if (!h) return ADRP_EINVAL;
    cudaStream_t s = (cudaStream_t)stream;
    return BY_PRECISION(h, do_reset<double>(typed<double>(h), env_mask, (double*)obs, s),
                        do_reset<float>(typed<float>(h), env_mask, (float*)obs, s));
}

int adrp_set_state(adrp_handle* h, const void* state, void* stream)
{
    if (!h || !state) return ADRP_EINVAL;
    cudaStream_t s = (cudaStream_t)stream;
    if (base(h)->precision == ADRP_F64) {
        auto* t = typed<double>(h);
        set_state_kernel<double><<<grid_for(t->P.Dpad), kBlock, 0, s>>>(t->P, t->B, (const double*)state);
    } else {
        auto* t = typed<float>(h);
        set_state_kernel<float><<<grid_for(t->P.Dpad), kBlock, 0, s>>>(t->P, t->B, (const float*)state);
    }
    base(h)->launches++;
    CUDA_TRY(base(h), cudaGetLastError());
    return ADRP_OK;
}

int adrp_get_state(adrp_handle* h, void* state, void* stream)
{
    if (!h || !state) return ADRP_EINVAL;
    cudaStream_t s = (cudaStream_t)stream;
    if (base(h)->precision == ADRP_F64) {
        auto* t = typed<double>(h);
        get_state_kernel<double><<<grid_for(t->P.Dpad), kBlock, 0, s>>>(t->P, t->B, (double*)state);
    } else {
        auto* t = typed<float>(h);
        get_state_kernel<float><<<grid_for(t->P.Dpad), kBlock, 0, s>>>(t->P, t->B, (float*)state);
    }
    base(h)->launches++;
    CUDA_TRY(base(h), cudaGetLastError());
    return ADRP_OK;
}

int adrp_step(adrp_handle* h, const void* action, const void* noise, void* obs, void* reward,
              uint8_t* terminated, uint8_t* truncated, uint8_t* drone_flags, void* stream)
{
    (void)noise;
    if (!h || !action || !obs || !reward || !terminated || !truncated) {
        if (h) base(h)->err = "adrp_step: NULL buffer";
        return ADRP_EINVAL;
    }
    cudaStream_t s = (cudaStream_t)stream;
    return BY_PRECISION(h, do_step<double>(typed<double>(h), (const double*)action, (double*)obs, (double*)reward, terminated, truncated, drone_flags, s),
                        do_step<float>(typed<float>(h), (const float*)action, (float*)obs, (float*)reward, terminated, truncated, drone_flags, s));
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

// Host-buffer step: pageable user memory -> pinned staging -> device -> kernel -> device -> pinned -> user.
template <typename T>
int do_step_host(Handle<T>* h, const void* action, void* obs, void* reward, uint8_t* te, uint8_t* tr, uint8_t* fl)
{
    int rc = ensure_staging(h);
    if (rc != ADRP_OK) return rc;
    const KParams<T>& P = h->P;
    const size_t nd = (size_t)P.N * P.M;
    cudaStream_t s = h->own_stream;
    memcpy(h->h_action, action, sizeof(T) * nd * P.A);
    CUDA_TRY(h, cudaMemcpyAsync(h->d_action, h->h_action, sizeof(T) * nd * P.A, cudaMemcpyHostToDevice, s));
    rc = do_step<T>(h, (const T*)h->d_action, (T*)h->d_obs, (T*)h->d_reward, h->d_term, h->d_trunc, h->d_flags, s);
    if (rc != ADRP_OK) return rc;
    CUDA_TRY(h, cudaMemcpyAsync(h->h_obs, h->d_obs, sizeof(T) * nd * P.O, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(h, cudaMemcpyAsync(h->h_reward, h->d_reward, sizeof(T) * P.N, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(h, cudaMemcpyAsync(h->h_term, h->d_term, P.N, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(h, cudaMemcpyAsync(h->h_trunc, h->d_trunc, P.N, cudaMemcpyDeviceToHost, s));
    if (fl) CUDA_TRY(h, cudaMemcpyAsync(h->h_flags, h->d_flags, nd, cudaMemcpyDeviceToHost, s));
    CUDA_TRY(h, cudaStreamSynchronize(s));
    memcpy(obs, h->h_obs, sizeof(T) * nd * P.O);
    memcpy(reward, h->h_reward, sizeof(T) * P.N);
    memcpy(te, h->h_term, P.N);
    memcpy(tr, h->h_trunc, P.N);
    if (fl) memcpy(fl, h->h_flags, nd);
    return ADRP_OK;
}

// DSLPIDControl.computeControl for n independent rows (control/DSLPIDControl.py:82-145).
template <typename T>
__global__ void dslpid_rows_kernel(T pid_gravity, T pid_kf, int model, int n, T dt, const T* pos, const T* quat,
                                   const T* vel, const T* tpos, const T* trpy, const T* tvel, const T* trates,
                                   T* ctrl, T* rpm, T* pos_e, T* yaw_e)
{
    const int r = blockIdx.x * blockDim.x + threadIdx.x;
    if (r >= n) return;
    PidState<T> c;
    T* cs = ctrl + 9 * r;
    c.ipos = {cs[0], cs[1], cs[2]}; c.irpy = {cs[3], cs[4], cs[5]}; c.last_rpy = {cs[6], cs[7], cs[8]};
    const V3<T> zero = {T(0), T(0), T(0)};
    auto v3 = [&](const T* p) -> V3<T> { return p ? V3<T>{p[3 * r], p[3 * r + 1], p[3 * r + 2]} : zero; };
    const Q4<T> q = {quat[4 * r], quat[4 * r + 1], quat[4 * r + 2], quat[4 * r + 3]};
    T out[4], ye;
    V3<T> pe;
    dslpid_control(pid_gravity, pid_kf, model, dt, v3(pos), q, v3(vel), v3(tpos), v3(trpy), v3(tvel), v3(trates), c, out, pe, ye);
    for (int k = 0; k < 4; ++k) rpm[4 * r + k] = out[k];
    if (pos_e) { pos_e[3 * r] = pe.x; pos_e[3 * r + 1] = pe.y; pos_e[3 * r + 2] = pe.z; }
    if (yaw_e) yaw_e[r] = ye;
    cs[0] = c.ipos.x; cs[1] = c.ipos.y; cs[2] = c.ipos.z; cs[3] = c.irpy.x; cs[4] = c.irpy.y; cs[5] = c.irpy.z;
    cs[6] = c.last_rpy.x; cs[7] = c.last_rpy.y; cs[8] = c.last_rpy.z;
}

}  // namespace

extern "C" {

int adrp_step_host(adrp_handle* h, const void* action, void* obs, void* reward, uint8_t* terminated,
                   uint8_t* truncated, uint8_t* drone_flags)
{
    if (!h || !action || !obs || !reward || !terminated || !truncated) return ADRP_EINVAL;
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
    cudaStream_t s = (cudaStream_t)stream;
    const unsigned grid = (unsigned)((n + 127) / 128);
    if (precision == ADRP_F64)
        dslpid_rows_kernel<double><<<grid, 128, 0, s>>>(g * drone->mass, drone->kf, drone_model, n, control_timestep,
            (const double*)pos, (const double*)quat, (const double*)vel, (const double*)target_pos, (const double*)target_rpy,
            (const double*)target_vel, (const double*)target_rates, (double*)ctrl_state, (double*)rpm, (double*)pos_e, (double*)yaw_e);
    else
        dslpid_rows_kernel<float><<<grid, 128, 0, s>>>((float)(g * drone->mass), (float)drone->kf, drone_model, n, (float)control_timestep,
            (const float*)pos, (const float*)quat, (const float*)vel, (const float*)target_pos, (const float*)target_rpy,
            (const float*)target_vel, (const float*)target_rates, (float*)ctrl_state, (float*)rpm, (float*)pos_e, (float*)yaw_e);
    return cudaGetLastError() == cudaSuccess ? ADRP_OK : ADRP_ECUDA;
}

int64_t adrp_launch_count(const adrp_handle* h) { return h ? base(h)->launches : 0; }

int adrp_set_profiling(adrp_handle* h, int enabled)
{
    if (!h) return ADRP_EINVAL;
    HandleBase* b = base(h);
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

const char* adrp_last_error(const adrp_handle* h) { return h ? base(h)->err.c_str() : g_create_error.c_str(); }

}  // extern "C"
