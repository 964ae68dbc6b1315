/* adrp.h -- C ABI of the B200-native batched drone-physics step (libadrp_b200.so).
 *
 * The reference (FelixWaiblinger/gym-pybullet-adrp) is pure Python and has no FFI of its own; its
 * extension points for this path are Python-level.  Each entry point below names the
 * reference interface it replaces.  Plain pointers and sizes only; every buffer passed to a
 * *_step / *_state call is caller-owned DEVICE memory unless the name ends in _host.  All functions
 * return 0 on success or a negative ADRP_E* code; adrp_last_error() gives the message.  A handle is
 * bound to one CUDA device and must be used from one host thread at a time; work is enqueued on the
 * caller's stream (a cudaStream_t passed as void*) without internal synchronisation, except the
 * *_host calls, which synchronise before returning.
 */
#ifndef ADRP_H
#define ADRP_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define ADRP_ABI_VERSION 6
#define ADRP_MAX_DRONES 32      /* drones per env: one env lives inside one warp */
#define ADRP_MAX_GATES 8
#define ADRP_MAX_OBSTACLES 8
#define ADRP_STATE_DIM 20       /* [pos3 quat4(xyzw) vel3 ang_v3(world) rpy_rates3(DYN body rates) last_rpm4] */

enum { ADRP_OK = 0, ADRP_EINVAL = -1, ADRP_ECUDA = -2, ADRP_EABI = -3, ADRP_EUNSUPPORTED = -4 };

enum adrp_precision { ADRP_F32 = 0, ADRP_F64 = 1 };
enum adrp_drone_model { ADRP_CF2X = 0, ADRP_CF2P = 1, ADRP_RACE = 2 };   /* utils/enums.py:8-14 */

/* Physics (utils/enums.py:18-26) as a bit mask.  PYB = 0, DYN = ADRP_PHY_DYN, PYB_GND = GND, ...
 * PYB_GND_DRAG_DW = GND|DRAG|DW; DYN|GND|DRAG|DW is the new combination of SURVEY.md F6. */
enum {
    ADRP_PHY_DYN = 1,     /* explicit model BaseAviary._dynamics (:822-896) instead of the Bullet free body */
    ADRP_PHY_GND = 2,     /* BaseAviary._groundEffect (:722-757) */
    ADRP_PHY_DRAG = 4,    /* BaseAviary._drag (:761-788) */
    ADRP_PHY_DW = 8,      /* BaseAviary._downwash (:792-818) */
    ADRP_PHY_GROUND = 16  /* declared plane contact response for the Bullet free body (PYB modes) */
};

enum adrp_action_type {          /* utils/enums.py:40-47 + CtrlAviary */
    ADRP_ACT_RPM_RAW = 0,        /* CtrlAviary._preprocessAction: clip(a, 0, MAX_RPM)  (CtrlAviary.py:140) */
    ADRP_ACT_RPM = 1,            /* HOVER_RPM*(1+0.05a)                    (BaseRLAviary.py:192) */
    ADRP_ACT_ONE_D_RPM = 2,      /* repeat(HOVER_RPM*(1+0.05a), 4)         (BaseRLAviary.py:224-225) */
    ADRP_ACT_PID = 3,            /* waypoint -> DSLPIDControl              (BaseRLAviary.py:193-207) */
    ADRP_ACT_ONE_D_PID = 4,      /* pos + 0.1*[0,0,a] -> DSLPIDControl     (BaseRLAviary.py:226-235) */
    ADRP_ACT_VEL = 5,            /* velocity command -> DSLPIDControl      (BaseRLAviary.py:208-223) */
    ADRP_ACT_MEL = 6,            /* [x,y,z,yaw] full-state setpoint -> Mellinger per substep (MultiRaceAviary.py:190-254) */
    ADRP_ACT_PID_FULL = 7,       /* [target_pos3, target_rpy3, target_vel3, target_rates3] -> DSLPIDControl.computeControl
                                    (control/DSLPIDControl.py:82-145), the examples/pid.py:138-147 driver loop fused */
    ADRP_ACT_MEL_FULL = 8,       /* (Command.FULLSTATE, args) tuples (MultiRaceAviary.py:198-205; MellingerControl.sendFullStateCmd,
                                    :510-543): [pos3, vel3, acc3, yaw, rpy_rates3 (rad/s)] -> Mellinger per substep */
    ADRP_ACT_MEL_CMD = 9         /* any (Command, args) tuple (utils/enums.py:58-70; low_level_control,
                                    MellingerControl.py:32-57): 16-float rows [adrp_command, args..., 0 padding] -- the
                                    high-level commander (take-off / land / go-to planner, :547-699) runs on the device */
};

/* Command (utils/enums.py:58-70) as the first float of an ADRP_ACT_MEL_CMD row, followed by the reference's args:
 *   FULLSTATE  pos3, vel3, acc3, yaw, rpy_rates3, t        TAKEOFF / LAND        height, duration
 *   TAKEOFFYAW / LANDYAW  height, duration, yaw            TAKEOFFVEL / LANDVEL  height, velocity, relative
 *   GOTO       pos3, yaw, duration, relative               STOP / NOTIFY         t          NONE  -            */
enum adrp_command {
    ADRP_CMD_NONE = 0, ADRP_CMD_FULLSTATE = 1, ADRP_CMD_TAKEOFF = 2, ADRP_CMD_TAKEOFFYAW = 3, ADRP_CMD_TAKEOFFVEL = 4,
    ADRP_CMD_LAND = 5, ADRP_CMD_LANDYAW = 6, ADRP_CMD_LANDVEL = 7, ADRP_CMD_STOP = 8, ADRP_CMD_GOTO = 9, ADRP_CMD_NOTIFY = 10
};

enum adrp_env_kind {
    ADRP_ENV_CTRL = 0,           /* CtrlAviary: obs (N,M,20), reward -1, flags false   (CtrlAviary.py:106-200) */
    ADRP_ENV_HOVER = 1,          /* HoverAviary (HoverAviary.py:68-117), obs (N,1,12+B*A) */
    ADRP_ENV_MULTIHOVER = 2,     /* MultiHoverAviary (MultiHoverAviary.py:71-130) */
    ADRP_ENV_RACE = 3            /* MultiRaceAviary (MultiRaceAviary.py:171-709), obs (N,M,49+6(M-1)) */
};

enum adrp_race_mode { ADRP_RACE_COMPARE = 0, ADRP_RACE_COMPETE = 1 };     /* utils/enums.py:84-87 */

/* per-drone flag bits written by adrp_step (MultiRaceAviary._computeTerminated, :674-698) */
enum {
    ADRP_FLAG_ELIMINATED = 1, ADRP_FLAG_FINISHED = 2, ADRP_FLAG_CRASHED = 4,
    ADRP_FLAG_OUT_OF_BOUNDS = 8, ADRP_FLAG_UNSTABLE = 16, ADRP_FLAG_TUMBLE_STOP = 32
};

/* adrp_config.wrappers: gym_pybullet_adrp/utils/wrapper.py fused into the race step as epilogues.
 *   ADRP_WRAP_OBS     DroneObservationWrapper.step (:39-63): the yaw component of every setpoint action is zeroed
 *                     and `terminated` is also raised once drone 0 has passed two gates.
 *   ADRP_WRAP_REWARD  RewardWrapper (:66-186): `reward` becomes the gate-progress shaping of drone 0
 *                     (xy progress + z progress towards the current gate, +5 per gate passed, -1 / +10 when the
 *                     env terminates without / with drone 0 having finished), with the wrapper's state
 *                     (current gate id, current target, previous position) kept per env on the device.
 * Composition is DroneObservationWrapper(RewardWrapper(env)): the reward sees the env's own `terminated`. */
enum { ADRP_WRAP_OBS = 1, ADRP_WRAP_REWARD = 2 };

/* URDF-level constants of one drone model (BaseAviary._parseURDFParameters, :989-1021). */
typedef struct adrp_drone_params {
    double mass, arm, kf, km, thrust2weight, max_speed_kmh;
    double gnd_eff_coeff, prop_radius, drag_coeff_xy, drag_coeff_z, dw_coeff_1, dw_coeff_2, dw_coeff_3;
    double ixx, iyy, izz;
    double collision_r, collision_h, collision_z_offset;
    double prop_x[4], prop_y[4];          /* prop link inertial origins: wrench arms in PYB modes (_physics, :706-712) */
    double pyb_mass, pyb_ixx, pyb_iyy, pyb_izz;  /* Bullet body mass/inertia (race: legacy cf2x.urdf) */
    double lin_damping, ang_damping;      /* btMultiBody defaults 0.04 (not disabled, BaseAviary.py:496-498) */
} adrp_drone_params;

/* level*.yaml track (config/level3.yaml:16-58) */
typedef struct adrp_track {
    int32_t num_gates, num_obstacles;
    double gates[ADRP_MAX_GATES][4];          /* nominal x, y, z, yaw */
    int32_t gate_type[ADRP_MAX_GATES];        /* 0 tall (portal.urdf), 1 low (low_portal.urdf) */
    double obstacles[ADRP_MAX_OBSTACLES][3];  /* nominal x, y, z */
    double bounds_hi[3];                      /* bounds[1]; only the upper row is used (:684) */
    double contact_margin;                    /* distance below which Bullet reports a contact point (0.02) */
    double visibility_range;                  /* constants.py:25 (0.45) */
} adrp_track;

/* level*.yaml randomisation and disturbances (config/level3.yaml:63-111) */
typedef struct adrp_randomization {
    int32_t random_gates_obstacles, random_drone_state, random_drone_inertia, disturbances;
    double gate_range[2], obstacle_range[2];
    double pos_range[3][2], rot_range[3][2];
    double mass_range[2], ixx_range[2], iyy_range[2], izz_range[2];
    double action_noise_std;                  /* N(0, std) newtons of thrust per motor per substep (:223-228) */
    double wind_lo[3], wind_hi[3];            /* U(lo, hi) world force per substep (:532-544) */
} adrp_randomization;

/* Everything BaseAviary.__init__ / MultiRaceAviary.__init__ take (BaseAviary.py:25-40,
 * MultiRaceAviary.py:31-44), flattened into one POD. */
typedef struct adrp_config {
    int32_t abi_version;          /* must be ADRP_ABI_VERSION */
    int32_t device;               /* CUDA device ordinal */
    int32_t precision;            /* adrp_precision: arithmetic AND buffer element type (float / double) */
    int32_t num_envs;             /* N */
    int32_t drones_per_env;       /* M = NUM_DRONES, 1..ADRP_MAX_DRONES */
    int32_t drone_model;          /* adrp_drone_model */
    int32_t physics;              /* ADRP_PHY_* mask */
    int32_t pyb_freq, ctrl_freq;  /* S = pyb_freq / ctrl_freq substeps are fused into one launch */
    int32_t action_type;          /* adrp_action_type */
    int32_t env_kind;             /* adrp_env_kind */
    int32_t race_mode;            /* adrp_race_mode */
    int32_t auto_reset;           /* 1: envs whose terminated|truncated is set are re-initialised inside the step */
    int32_t dslpid_model;         /* drone model the DSLPID mixer/constants come from (BaseRLAviary.py:75-76 always CF2X) */
    uint32_t bug_compat;          /* reserved: SURVEY Appendix E quirks, all reproduced (must be 0xFFFFFFFF) */
    int32_t wrappers;             /* ADRP_WRAP_* mask: the reference's gymnasium wrappers fused into the step (race env) */
    double g;                     /* 9.8 (BaseAviary.py:75) */
    double episode_len_sec;       /* HoverAviary 8 s (:52), race config episode_len_sec */
    double mel_mass;              /* firmware g_vehicleMass (pycffirmware, declared 0.027) */
    uint64_t seed;                /* Philox key; subsequence = global env index + env_offset */
    uint64_t env_offset;          /* first global env index of this shard (multi-GPU sharding) */
    double init_xyz[ADRP_MAX_DRONES][3];   /* INIT_XYZS (BaseAviary.py:193-202) */
    double init_rpy[ADRP_MAX_DRONES][3];   /* INIT_RPYS */
    double init_vel[ADRP_MAX_DRONES][3];   /* race init_states.vel */
    double init_pqr[ADRP_MAX_DRONES][3];   /* race init_states.pqr */
    double target_pos[ADRP_MAX_DRONES][3]; /* Hover/MultiHover TARGET_POS */
    adrp_drone_params drone;
    adrp_drone_params dslpid_drone;        /* URDF constants the DSLPID controller reads (BaseControl.py:35-39) */
    adrp_track track;
    adrp_randomization rnd;
    /* CtrlAviary-kind envs: 0 or 20 = the reference's 20-float row of _getDroneStateVector (envs/CtrlAviary.py:106-117,
     * BaseAviary.py:545-565); 12 = only its kinematic columns [pos3, rpy3, vel3, ang_vel3] -- the twelve that
     * BaseRLAviary._computeObs picks from it (:310-317) -- for callers who do not read quaternion and last RPM: 40 % less
     * observation traffic to HBM and over PCIe.  Other env kinds: must be 0. */
    int32_t obs_columns;
    int32_t reserved0;            /* 0 */
} adrp_config;

typedef struct adrp_handle adrp_handle;

/* Replaces Env.__init__ (BaseAviary.py:25-218; MultiRaceAviary.py:31-123): allocates the SoA state of
 * N x M drones in HBM and puts every drone at its initial pose (no randomisation until adrp_reset). */
int adrp_create(const adrp_config* cfg, adrp_handle** out);
/* Replaces Env.close (BaseAviary.py:420-425). */
int adrp_destroy(adrp_handle* h);

int adrp_obs_dim(const adrp_handle* h);          /* floats per drone in `obs` */
int adrp_action_dim(const adrp_handle* h);       /* floats per drone in `action` */
int adrp_noise_dim(const adrp_handle* h);        /* floats per drone per substep in `noise` (7: wind3 + motor4) */
int adrp_reset_draws_dim(const adrp_handle* h);  /* floats per env in `draws`, see DESIGN.md */
int adrp_substeps(const adrp_handle* h);         /* S */

/* Replaces Env.reset (BaseAviary.py:223-258; MultiRaceAviary.py:127-167).  env_mask: N bytes (device)
 * or NULL = all envs.  draws: optional (N, reset_draws_dim) uniform[0,1) numbers (device) replacing
 * the in-kernel Philox stream -- the reference's RNG is unseeded, so parity tests inject
 * them.  obs: optional (N, M, obs_dim) initial observation. */
int adrp_reset(adrp_handle* h, const uint8_t* env_mask, const void* draws, void* obs, void* stream);

/* State injection / read-back, (N, M, ADRP_STATE_DIM) in the handle's precision (device).
 * Stand in for p.resetBasePositionAndOrientation / resetBaseVelocity and
 * BaseAviary._getDroneStateVector (:545-565). */
int adrp_set_state(adrp_handle* h, const void* state, void* stream);
int adrp_get_state(adrp_handle* h, void* state, void* stream);

/* Replaces Env.step (BaseAviary.py:262-387; MultiRaceAviary.py:171-270): ONE kernel launch advances
 * every drone by S physics substeps (controller, force models, integrator, contact / gate tests) and
 * writes observation, reward and flags.
 *   action      (N, M, action_dim)            handle precision
 *   noise       (N, M, S, noise_dim) or NULL  injected disturbance draws (race env), handle precision
 *   obs         (N, M, obs_dim)               handle precision
 *   reward      (N)                           handle precision
 *   terminated, truncated (N) bytes; drone_flags (N, M) bytes (may be NULL) */
int adrp_step(adrp_handle* h, const void* action, const void* noise, void* obs, void* reward,
              uint8_t* terminated, uint8_t* truncated, uint8_t* drone_flags, void* stream);

/* n_steps consecutive Env.step calls in ONE launch: the rigid-body and controller state of every drone stays in
 * registers from one control step to the next, and per-launch costs (launch latency, state load / store, block
 * scheduling) are paid once per n_steps -- what a scripted rollout (examples/pid.py:138-147: a fixed DSLPID target;
 * user_controller/HardCodedController.py:166-176: a precomputed setpoint sequence) or a random-action rollout
 * (examples/learn.py's shape) needs.  Results equal those of n_steps adrp_step calls: bit for bit in the fp64 build, to
 * rounding in the fp32 build (its n-step kernels are separate instances, so that the single-step ones keep their
 * register count; nvcc contracts a few multiply-adds differently); flags, counters and reset draws are the same.
 *   actions_per_step  ADRP_ACTIONS_PER_STEP: action is (n_steps, N, M, action_dim) and noise (n_steps, N, M, S, noise_dim);
 *                     ADRP_ACTIONS_HELD: one (N, M, action_dim) slice (and one noise slice) is held for every step;
 *                     ADRP_ACTIONS_PHILOX: `action` is ignored (may be NULL) and every action component is drawn in the
 *                       kernel, a ~ U[-1, 1) -- what action_space.sample() of a random policy draws over the reference's
 *                       Box(-1, 1) spaces (envs/BaseRLAviary.py:129-135, envs/MultiRaceAviary.py:284-287), for action types
 *                       of at most 4 components: a = ((x >> 8) - 2^23) * 2^-23 with x word k of Philox4x32-10(key =
 *                       adrp_config.seed; counter = (env_offset + env, 192 + drone, rollout step, 0)), the same in both
 *                       precisions and on any sharding.  The rollout step counts the control steps drawn on this handle so
 *                       far (0 at adrp_create, advanced by n_steps per call; adrp_set_rollout_step).  noise: per step.
 *   outputs_per_step  != 0: obs (n_steps, N, M, obs_dim), reward / terminated / truncated (n_steps, N),
 *                           drone_flags (n_steps, N, M): every step's outputs are kept;
 *                     == 0: the same (N, M, obs_dim) ... buffers are overwritten every step (the last step's remain) */
enum { ADRP_ACTIONS_HELD = 0, ADRP_ACTIONS_PER_STEP = 1, ADRP_ACTIONS_PHILOX = 2 };
int adrp_step_n(adrp_handle* h, int n_steps, int actions_per_step, int outputs_per_step, const void* action,
                const void* noise, void* obs, void* reward, uint8_t* terminated, uint8_t* truncated,
                uint8_t* drone_flags, void* stream);

/* The rollout step of ADRP_ACTIONS_PHILOX (the third Philox counter word of the next draw): set it to replay a rollout or
 * to continue one from a checkpoint.  No counterpart in the reference, whose action_space.sample() is unseeded. */
int adrp_set_rollout_step(adrp_handle* h, uint32_t step);

/* Same call with HOST buffers (the path a NumPy user of the reference takes): copies the action to the
 * device, launches, copies obs / reward / flags back and synchronises. */
int adrp_step_host(adrp_handle* h, const void* action, void* obs, void* reward,
                   uint8_t* terminated, uint8_t* truncated, uint8_t* drone_flags);

/* Replaces DSLPIDControl.computeControl (control/DSLPIDControl.py:82-145) for n drones at once
 * (device buffers, SoA-free plain rows): pos (n,3) quat (n,4) vel (n,3) target_pos/rpy/vel/rates (n,3),
 * ctrl_state (n,9) in/out = [integral_pos_e3, integral_rpy_e3, last_rpy3]; outputs rpm (n,4),
 * pos_e (n,3), yaw_e (n). */
int adrp_dslpid_control(int precision, int drone_model, const adrp_drone_params* drone, double g, int n,
                        double control_timestep, const void* pos, const void* quat, const void* vel,
                        const void* target_pos, const void* target_rpy, const void* target_vel,
                        const void* target_rates, void* ctrl_state, void* rpm, void* pos_e, void* yaw_e,
                        void* stream);

/* Number of kernel launches issued through this handle since creation (bench.py's gpu_launches). */
int64_t adrp_launch_count(const adrp_handle* h);
/* Average device time (ms) of the step kernel over the launches since the last call with reset != 0,
 * measured with CUDA events on the launching stream when profiling is enabled. */
int adrp_set_profiling(adrp_handle* h, int enabled);
double adrp_step_kernel_ms(adrp_handle* h, int reset);

/* Measured FP32 pipe peak of `device` (TFLOP/s, FMA = 2 flop): a register-resident chain of independent FMAs,
 * timed with CUDA events; the denominator of bench.py's fp32 roofline (MEASURED_PEAKS.json has no FP32 figure).
 * Returns a negative value on error. */
double adrp_fp32_peak_tflops(int device);

/* NO COUNTERPART IN THE REFERENCE: added for callers that cannot call cudaMalloc
 * (code under CUDA graph capture, tight serving loops, harnesses that forbid it).
 *
 * One control step of Env.step on caller-owned rows, WITHOUT a handle and WITHOUT any allocation: the structure-of-arrays
 * working copy of the state that step_kernel runs on (what adrp_create cudaMallocs for a handle) lives in a file-scope
 * `static __device__` array of ADRP_STATIC_SCRATCH_BYTES inside the library, which every GPU that loads the library holds
 * until the process exits.  The call enqueues, on `stream`: rows -> SoA scratch, the step kernel (the same instances adrp_step
 * launches, so results are bit-identical to adrp_set_state + adrp_step + adrp_get_state on a handle of the same config),
 * SoA scratch -> rows.  It calls no cudaMalloc / cudaFree / cudaMallocHost and no allocating library routine.
 *   cfg         env_kind must be ADRP_ENV_CTRL and action_type ADRP_ACT_RPM_RAW (any physics mask, drone model, M <= 32,
 *               either precision); num_envs * lanes-per-env (M rounded up to a power of two) * 20 * sizeof(element)
 *               + 4 * num_envs must not exceed ADRP_STATIC_SCRATCH_BYTES (fp32: about 100 000 padded drones), else
 *               ADRP_EINVAL with a message naming the need and the limit (adrp_last_error(NULL))
 *   state_in    (N, M, ADRP_STATE_DIM) device, cfg->precision        action  (N, M, 4) device
 *   state_out   (N, M, ADRP_STATE_DIM) device (may alias state_in)   obs     (N, M, 20) device
 *   reward      (N) device; terminated, truncated (N) bytes device
 * The scratch is SHARED by every call in the process and keeps the last call's contents: the call overwrites all of it that
 * it reads (the per-env step counters are cleared to zero first), so successive calls are independent of each other, but
 * two calls that overlap on the GPU (different streams, or graphs replayed concurrently) RACE on it -- serialise them.
 * First use on a device resolves the array's device address (cudaGetSymbolAddress) and loads the kernels: make one warm-up
 * call outside a graph capture. */
#define ADRP_STATIC_SCRATCH_BYTES (8u << 20)
int adrp_step_rows_static(const adrp_config* cfg, const void* state_in, const void* action, void* state_out, void* obs,
                          void* reward, uint8_t* terminated, uint8_t* truncated, void* stream);

const char* adrp_last_error(const adrp_handle* h);   /* h may be NULL: error of the last failed adrp_create */
int adrp_abi_version(void);
int adrp_config_size(void);                           /* sizeof(adrp_config) for binding sanity checks */

#ifdef __cplusplus
}
#endif
#endif /* ADRP_H */
