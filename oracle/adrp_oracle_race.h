/* TEST INFRASTRUCTURE -- race env + Mellinger restatement (REAL-generic).  Filled in below. */
typedef struct FN(orc_race) { int dummy; } FN(orc_race);
static FN(orc_race)* FN(race_create)(FN(orc)* h) { (void)h; return NULL; }
static void FN(race_destroy)(FN(orc_race)* r) { (void)r; }
static int FN(race_draws_dim)(const FN(orc)* h) { (void)h; return 0; }
static void FN(race_init_env)(FN(orc)* h, int n, const REAL* draws) { (void)h; (void)n; (void)draws; }
static void FN(race_write_obs)(FN(orc)* h, int n, REAL* obs) { (void)h; (void)n; (void)obs; }
static void FN(race_step_env)(FN(orc)* h, int n, const REAL* action, const REAL* noise, REAL* obs, REAL* reward,
                              uint8_t* te, uint8_t* tr, uint8_t* fl)
{ (void)h; (void)n; (void)action; (void)noise; (void)obs; (void)reward; (void)te; (void)tr; (void)fl; }
