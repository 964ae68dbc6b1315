/* TEST INFRASTRUCTURE -- mechanical operation count of the CPU oracle ("recount mechanically from the
 * frozen oracle").
 *
 * The oracle headers are compiled once more, as C++, with REAL and the firmware's `float` replaced by counting
 * wrappers: every +, -, *, comparison, division, sqrt and libm call the restatement executes increments a counter.
 * Same statements, same control flow, same results as oracle/liboracle.so (fp64 build) -- only instrumented.
 * Conventions: add / sub / mul / compare = 1 flop each (no FMA in the restatement: a multiply-add is 2);
 * div, sqrt, sin, cos, tan, atan2, asin, exp, log are "SFU" operations, counted separately; negation, fabs,
 * conversions and copies are free.  Driver: oracle/flopcount.py.  Build: make -C oracle liboracle_count.so
 * The product never links or loads this file.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

#include <type_traits>

#include "../include/adrp.h"

namespace cnt {
enum { ADD = 0, MUL, CMP, DIV, SQRT, SINCOS, TAN, ATAN2, ASIN, EXP, LOG, NKIND };
static unsigned long long c[NKIND];
}  // namespace cnt

template <typename B>
struct Counted {
    B v;
    Counted() = default;
    template <typename U, typename = typename std::enable_if<std::is_arithmetic<U>::value>::type>
    Counted(U x) : v((B)x) {}
    /* widening float -> double is implicit like in C; narrowing needs the cast the oracle already writes */
    template <typename B2, typename = typename std::enable_if<(sizeof(B2) < sizeof(B))>::type>
    Counted(Counted<B2> o) : v((B)o.v) {}
    template <typename B2, typename = typename std::enable_if<(sizeof(B2) > sizeof(B))>::type, typename = void>
    explicit Counted(Counted<B2> o) : v((B)o.v) {}
    template <typename U, typename = typename std::enable_if<std::is_arithmetic<U>::value>::type>
    explicit operator U() const { return (U)v; }

    friend Counted operator+(Counted a, Counted b) { cnt::c[cnt::ADD]++; return Counted(a.v + b.v); }
    friend Counted operator-(Counted a, Counted b) { cnt::c[cnt::ADD]++; return Counted(a.v - b.v); }
    friend Counted operator*(Counted a, Counted b) { cnt::c[cnt::MUL]++; return Counted(a.v * b.v); }
    friend Counted operator/(Counted a, Counted b) { cnt::c[cnt::DIV]++; return Counted(a.v / b.v); }
    friend Counted operator-(Counted a) { return Counted(-a.v); }
    friend Counted operator+(Counted a) { return a; }
    Counted& operator+=(Counted b) { cnt::c[cnt::ADD]++; v += b.v; return *this; }
    Counted& operator-=(Counted b) { cnt::c[cnt::ADD]++; v -= b.v; return *this; }
    Counted& operator*=(Counted b) { cnt::c[cnt::MUL]++; v *= b.v; return *this; }
    Counted& operator/=(Counted b) { cnt::c[cnt::DIV]++; v /= b.v; return *this; }
    friend bool operator<(Counted a, Counted b) { cnt::c[cnt::CMP]++; return a.v < b.v; }
    friend bool operator>(Counted a, Counted b) { cnt::c[cnt::CMP]++; return a.v > b.v; }
    friend bool operator<=(Counted a, Counted b) { cnt::c[cnt::CMP]++; return a.v <= b.v; }
    friend bool operator>=(Counted a, Counted b) { cnt::c[cnt::CMP]++; return a.v >= b.v; }
    friend bool operator==(Counted a, Counted b) { cnt::c[cnt::CMP]++; return a.v == b.v; }
    friend bool operator!=(Counted a, Counted b) { cnt::c[cnt::CMP]++; return a.v != b.v; }
};
typedef Counted<double> CD;
typedef Counted<float> CF;

#define CNT_FN1(name, kind, fd, ff)                                                \
    static inline CD name(CD x) { cnt::c[cnt::kind]++; return CD(fd(x.v)); }       \
    static inline CF name(CF x) { cnt::c[cnt::kind]++; return CF(ff(x.v)); }
CNT_FN1(c_sin, SINCOS, ::sin, ::sinf)
CNT_FN1(c_cos, SINCOS, ::cos, ::cosf)
CNT_FN1(c_tan, TAN, ::tan, ::tanf)
CNT_FN1(c_sqrt, SQRT, ::sqrt, ::sqrtf)
CNT_FN1(c_asin, ASIN, ::asin, ::asinf)
CNT_FN1(c_exp, EXP, ::exp, ::expf)
CNT_FN1(c_log, LOG, ::log, ::logf)
static inline CD c_fabs(CD x) { return CD(::fabs(x.v)); }
static inline CF c_fabs(CF x) { return CF(::fabsf(x.v)); }
static inline CD c_atan2(CD y, CD x) { cnt::c[cnt::ATAN2]++; return CD(::atan2(y.v, x.v)); }
static inline CF c_atan2(CF y, CF x) { cnt::c[cnt::ATAN2]++; return CF(::atan2f(y.v, x.v)); }
/* float libm names the firmware restatement calls directly */
static inline CF sqrtf(CF x) { return c_sqrt(x); }
static inline CF tanf(CF x) { return c_tan(x); }
static inline CF cosf(CF x) { return c_cos(x); }
static inline CF sinf(CF x) { return c_sin(x); }
static inline CF fabsf(CF x) { return c_fabs(x); }
static inline CF copysignf(CF a, CF b) { return CF(::copysignf(a.v, b.v)); }      /* high-level commander (adrp_oracle_planner.h) */
static inline CF fmodf(CF a, CF b) { cnt::c[cnt::DIV]++; return CF(::fmodf(a.v, b.v)); }
static inline CF floorf(CF a) { return CF(::floorf(a.v)); }
static inline bool isfinite(CF x) { return ::isfinite(x.v); }
static inline bool isfinite(CD x) { return ::isfinite(x.v); }

#define REAL CD
#define SFX f64
#define SIN c_sin
#define COS c_cos
#define SQRT c_sqrt
#define FABS c_fabs
#define ATAN2 c_atan2
#define ASIN c_asin
#define EXP c_exp
#define LOG c_log
#define TAN c_tan
#define FW_ATAN2F(y, x) ((CF)c_atan2((CD)(y), (CD)(x)))
#define FW_SINF(x) ((CF)c_sin((CD)(x)))
#define FW_COSF(x) ((CF)c_cos((CD)(x)))
#define float CF                      /* the firmware core's arithmetic type, this translation unit only */
#include "adrp_oracle_impl.h"
#include "adrp_oracle_race.h"
#include "adrp_oracle_api.h"
#undef float

extern "C" {

int cnt_create(const adrp_config* cfg, void** out)
{
    if (!cfg || cfg->abi_version != ADRP_ABI_VERSION) return ADRP_EABI;
    *out = (void*)create_f64(cfg);
    return *out ? ADRP_OK : ADRP_EINVAL;
}
int cnt_destroy(void* h) { destroy_f64((orc_f64*)h); return ADRP_OK; }
int cnt_obs_dim(void* h) { return ((orc_f64*)h)->O; }
int cnt_action_dim(void* h) { return ((orc_f64*)h)->A; }
int cnt_reset(void* h, const double* draws, double* obs)
{
    reset_f64((orc_f64*)h, NULL, (const CD*)draws, (CD*)obs);
    return ADRP_OK;
}
int cnt_set_state(void* h, const double* state)
{
    orc_f64* o = (orc_f64*)h;
    memcpy(o->state, state, sizeof(double) * 20 * (size_t)o->N * o->M);
    return ADRP_OK;
}
int cnt_get_state(void* h, double* state)
{
    orc_f64* o = (orc_f64*)h;
    memcpy(state, o->state, sizeof(double) * 20 * (size_t)o->N * o->M);
    return ADRP_OK;
}
int cnt_step(void* h, const double* action, const double* noise, double* obs, double* reward, uint8_t* te, uint8_t* tr,
             uint8_t* fl)
{
    step_f64((orc_f64*)h, (const CD*)action, (const CD*)noise, (CD*)obs, (CD*)reward, te, tr, fl, 1);
    return ADRP_OK;
}
void cnt_counts(unsigned long long* out, int reset)
{
    for (int k = 0; k < cnt::NKIND; ++k) { out[k] = cnt::c[k]; if (reset) cnt::c[k] = 0; }
}
int cnt_kinds(void) { return cnt::NKIND; }

}  /* extern "C" */
