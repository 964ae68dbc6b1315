/* TEST INFRASTRUCTURE -- oracle handle management and the env loop (REAL-generic, see adrp_oracle.c). */

static FN(orc)* FN(create)(const adrp_config* c)
{
    if (c->drones_per_env < 1 || c->drones_per_env > ADRP_MAX_DRONES || c->num_envs < 1) return NULL;
    if (c->pyb_freq % c->ctrl_freq != 0) return NULL;                    /* BaseAviary.py:79-80 */
    FN(orc)* h = (FN(orc)*)calloc(1, sizeof(FN(orc)));
    h->cfg = *c;
    FN(derive)(c, &h->D);
    h->N = c->num_envs; h->M = c->drones_per_env; h->S = c->pyb_freq / c->ctrl_freq;
    h->A = FN(action_dim)(c->action_type);
    h->B = c->env_kind == ADRP_ENV_HOVER || c->env_kind == ADRP_ENV_MULTIHOVER ? c->ctrl_freq / 2 : 0;   /* BaseRLAviary.py:66 */
    if (c->env_kind == ADRP_ENV_CTRL) h->O = c->obs_columns == 12 ? 12 : 20;
    else if (c->env_kind == ADRP_ENV_RACE) h->O = 49 + (c->race_mode == ADRP_RACE_COMPETE ? 6 * (h->M - 1) : 0);
    else h->O = 12 + h->B * h->A;
    size_t D = (size_t)h->N * h->M;
    h->state = (REAL*)calloc(D * 20, sizeof(REAL));
    h->pid = (REAL*)calloc(D * 9, sizeof(REAL));
    h->ring = (REAL*)calloc(D * (size_t)(h->B * h->A) + 1, sizeof(REAL));
    h->step_counter = (int32_t*)calloc(h->N, sizeof(int32_t));
    if (c->env_kind == ADRP_ENV_RACE) h->race = FN(race_create)(h);
    for (int n = 0; n < h->N; ++n) {
        FN(init_env)(h, n);
        if (h->race) FN(race_init_env)(h, n, NULL);
    }
    return h;
}

static void FN(destroy)(FN(orc)* h)
{
    if (h->race) FN(race_destroy)(h->race);
    free(h->state); free(h->pid); free(h->ring); free(h->step_counter); free(h);
}

static int FN(reset_draws_dim)(const FN(orc)* h)
{
    return h->race ? FN(race_draws_dim)(h) : 0;
}

/* Env.reset (BaseAviary.py:223-258; MultiRaceAviary.py:127-167) */
static void FN(reset)(FN(orc)* h, const uint8_t* mask, const REAL* draws, REAL* obs)
{
    const int R = FN(reset_draws_dim)(h);
    for (int n = 0; n < h->N; ++n) {
        if (mask && !mask[n]) continue;
        FN(init_env)(h, n);
        if (h->race) FN(race_init_env)(h, n, draws ? draws + (size_t)n * R : NULL);
    }
    if (obs)
        for (int n = 0; n < h->N; ++n) {
            if (mask && !mask[n]) continue;
            if (!h->race) { FN(write_obs)(h, n, obs); continue; }
            /* MultiRaceAviary.reset (:143-144): initial_obs comes from BaseAviary.reset, BEFORE _drone_init
             * randomises the drones -- nominal INIT_XYZS / INIT_RPYS (= config rpy * pi/180, :95) at rest, seen against
             * the already randomised track. */
            REAL keep[ADRP_MAX_DRONES * 20];
            REAL* env = h->state + (size_t)n * h->M * 20;
            memcpy(keep, env, sizeof(REAL) * 20 * h->M);
            for (int i = 0; i < h->M; ++i) {
                REAL* st = env + 20 * i;
                REAL e[3];
                for (int k = 0; k < 20; ++k) st[k] = 0;
                for (int k = 0; k < 3; ++k) {
                    st[k] = (REAL)h->cfg.init_xyz[i][k];
                    e[k] = (REAL)(h->cfg.init_rpy[i][k] * (3.14159265358979323846 / 180.0));
                }
                FN(quat_from_euler)(e, st + 3);
                FN(pose_round_trip)(st + 3);
            }
            FN(race_write_obs)(h, n, obs);
            memcpy(env, keep, sizeof(REAL) * 20 * h->M);
        }
}

static void FN(step)(FN(orc)* h, const REAL* action, const REAL* noise, REAL* obs, REAL* reward,
                     uint8_t* terminated, uint8_t* truncated, uint8_t* flags, int nthreads)
{
    if (nthreads < 1) nthreads = 1;
#pragma omp parallel for schedule(static) num_threads(nthreads) if (nthreads > 1)
    for (int n = 0; n < h->N; ++n) {
        if (h->race) FN(race_step_env)(h, n, action, noise, obs, reward, terminated, truncated, flags);
        else FN(step_env)(h, n, action, obs, reward, terminated, truncated);
    }
}

static void FN(dslpid_rows)(int model, const adrp_drone_params* p, double g, int n, double dt,
                            const void* pos, const void* quat, const void* vel, const void* tpos, const void* trpy,
                            const void* tvel, const void* trates, void* ctrl, void* rpm, void* pos_e, void* yaw_e)
{
    FN(orc_drone) D;
    memset(&D, 0, sizeof(D));
    D.pid_GRAVITY = (REAL)(g * p->mass);
    D.pid_KF = (REAL)p->kf;
    for (int i = 0; i < n; ++i)
        FN(dslpid)(&D, model, (REAL)dt, (const REAL*)pos + 3 * i, (const REAL*)quat + 4 * i, (const REAL*)vel + 3 * i,
                   (const REAL*)tpos + 3 * i, (const REAL*)trpy + 3 * i, (const REAL*)tvel + 3 * i,
                   (const REAL*)trates + 3 * i, (REAL*)ctrl + 9 * i, (REAL*)rpm + 4 * i, (REAL*)pos_e + 3 * i,
                   (REAL*)yaw_e + i);
}
