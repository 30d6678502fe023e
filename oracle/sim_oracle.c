/*
 * oracle/sim_oracle.c -- CPU restatement of HDDM's Euler-Maruyama simulators.  TEST INFRASTRUCTURE ONLY
 * (same rules as wfpt_oracle.c: nothing under hddm_b200/ may load or call it).
 *
 * Follows /root/reference/src/cddm_data_simulation.pyx, float32 state and accumulated float32 time included:
 *   model 0 ddm :294-375            1 full_ddm_hddm_base :174-284     2 ddm_flexbound (constant bound) :454-553
 *   model 3 ddm_flexbound + angle bound, 4 + Weibull bound (graveyard/boundary_functions.py:18-28, :50-63; the
 *           boundary table b[ix] = a (+|*) f(t_s[ix]) is built per parameter set as the reference does, :491-509)
 *   model 5 full_ddm :747-869       6 ddm_sdv :876-994   7 ornstein_uhlenbeck :1119-1219   8 levy_flexbound :642-741
 *   model 9/10 ddm_flex :1001-1113 with gamma_drift (constant / angle bound), 11/12 with ds_conflict_drift
 *           (graveyard/drift_functions.py:11-103; the drift table drift[ix] = v + f(t_s[ix]) is built per set, :1054-1056)
 *   model 13 ddm_cov :384-449
 * Gaussian draws: Marsaglia polar method as in :48-60; alpha-stable draws: :28-40.  The RNG is xoshiro256** instead
 * of C rand(), so results agree with the reference in distribution, not draw by draw.
 *
 * Parity pinning: the reference's own tests pin only shapes and metadata keys for simulators
 * (hddm/tests/test_simulators.py:65-87) -> PARITY UNPINNED for simulator numerics.  tests/test_sim_oracle.py checks
 * this port distributionally (two-sample KS on RTs, choice-probability intervals) against the reference's Cython
 * compiled unmodified into oracle/_ref, and against the analytic WFPT for the constant-bound DDM.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif
#ifndef M_PI_2
#define M_PI_2 1.57079632679489661923
#endif

typedef struct { uint64_t s[4]; } rng_t;
static inline uint64_t rotl(uint64_t x, int k) { return (x << k) | (x >> (64 - k)); }
static uint64_t splitmix(uint64_t* x) { uint64_t z = (*x += 0x9e3779b97f4a7c15ULL); z = (z ^ (z >> 30)) * 0xbf58476d1ce4e5b9ULL; z = (z ^ (z >> 27)) * 0x94d049bb133111ebULL; return z ^ (z >> 31); }
static void rng_seed(rng_t* r, uint64_t seed) { for (int i = 0; i < 4; i++) r->s[i] = splitmix(&seed); }
static inline uint64_t rng_next(rng_t* r)
{
    const uint64_t res = rotl(r->s[1] * 5, 7) * 9, t = r->s[1] << 17;
    r->s[2] ^= r->s[0]; r->s[3] ^= r->s[1]; r->s[1] ^= r->s[2]; r->s[0] ^= r->s[3]; r->s[2] ^= t; r->s[3] = rotl(r->s[3], 45);
    return res;
}
/* :21-23 random_uniform: a float in [0, 1] */
static inline float random_uniform(rng_t* r) { return (float)((double)(rng_next(r) >> 11) * (1.0 / 9007199254740992.0)); }
/* :48-60 random_gaussian (polar method), float32 arithmetic */
static float random_gaussian(rng_t* r)
{
    float x1, x2, w = 2.0f;
    while (w >= 1.0f) {
        x1 = 2.0f * random_uniform(r) - 1.0f;
        x2 = 2.0f * random_uniform(r) - 1.0f;
        w = x1 * x1 + x2 * x2;
    }
    w = (float)pow((-2.0 * log(w)) / w, 0.5);
    return x1 * w;
}
/* :28-40 random_stable */
static float random_stable(rng_t* r, float alpha)
{
    float u = (float)(M_PI * (random_uniform(r) - 0.5));
    float w = (float)(-log(random_uniform(r)));
    if (alpha == 1.0f) return (float)((1.0 / M_PI_2) * (M_PI_2 * tan(u)));
    return (float)((sin(alpha * u) / (pow(cos(u), 1 / alpha))) * pow(cos(u - (alpha * u)) / w, (1.0 - alpha) / alpha));
}
static inline int sgn(float x) { return (x > 0) - (x < 0); }

int orc_sim_run(int model, const float* theta, int n_par, int n_sets, int64_t n_samples, float dt, float max_t, float s,
                uint64_t seed, float* rts, int32_t* choices)
{
    static const int npar[14] = {4, 7, 4, 5, 6, 7, 5, 5, 5, 7, 8, 10, 11, 4};
    if (model < 0 || model > 13 || n_par != npar[model]) return 1;
    rng_t R;
    rng_seed(&R, seed);
    const float sqrt_st = (float)sqrt(dt) * s;
    const int nb = (int)((max_t / dt) + 1) + 2;
    float* b = (float*)malloc(sizeof(float) * (size_t)(nb + 1));
    float* dr = (float*)malloc(sizeof(float) * (size_t)(nb + 1));
    if (!b || !dr) return 2;
    for (int k = 0; k < n_sets; k++) {
        const float* th = theta + (size_t)k * n_par;
        float v = th[0], a = th[1], z = th[2];
        float t = th[3], g = 0, alpha = 0, dt_alpha = 0, ang = (model == 3) ? th[4] : 0;
        if (model == 10) ang = th[4];
        if (model == 11 || model == 12) { v = 0; a = th[0]; z = th[1]; t = th[2]; if (model == 12) ang = th[10]; }
        if (model == 7) { g = th[3]; t = th[4]; }
        if (model == 8) { alpha = th[3]; t = th[4]; dt_alpha = (float)(s * pow(dt, 1.0 / alpha)); }
        const int zero_a = (model == 0 || model == 1 || model == 13);
        const int tdrift = (model >= 9 && model <= 12);
        /* boundary table on t_s = arange(0, max_t + dt, dt).astype(float32)  (:489-509) */
        for (int ix = 0; ix <= nb; ix++) {
            const float ts = (float)((double)ix * (double)dt);
            if (model == 3 || model == 10 || model == 12) b[ix] = a + ts * (-sinf(ang) / cosf(ang));
            else if (model == 4) b[ix] = (float)(a * exp(-pow((double)ts / th[5], (double)th[4])));
            else b[ix] = a;
            if (model == 9 || model == 10) {
                const int o = (model == 10) ? 5 : 4;
                const double shape = th[o], scale = th[o + 1], cc = th[o + 2];
                const double num = pow((double)ts, shape - 1) * exp(-(double)ts / scale);
                const double div = pow(shape - 1, shape - 1) * pow(scale, shape - 1) * exp(-(shape - 1));
                dr[ix] = (float)(v + cc * (num / div));
            } else if (model == 11 || model == 12) {
                const double wt = ((double)th[3] - th[7]) * exp(-((double)th[5] * ts)) + th[7];
                const double wd = ((double)th[4] - 0) * exp(-((double)th[6] * ts)) + 0;
                dr[ix] = (float)(v + ((wt * th[8]) + (wd * th[9])));
            } else dr[ix] = v;
        }
        for (int64_t n = 0; n < n_samples; n++) {
            float y = zero_a ? z * a : (-1) * b[0] + (z * 2 * b[0]);
            float drift = v * dt, t_nd = t;
            if (model == 1 || model == 5) {
                const float sz = th[4], sv = th[5], st = th[6];
                drift = (v + sv * random_gaussian(&R)) * dt;
                t_nd = t + (2 * (random_uniform(&R) - 0.5f) * st);
                y += 2 * (random_uniform(&R) - 0.5f) * sz;
            }
            if (model == 6) drift = (v + th[4] * random_gaussian(&R)) * dt;
            float tp = 0.0f;
            int ix = 0;
            while ((zero_a ? (y >= 0 && y <= a) : (y >= (-1) * b[ix] && y <= b[ix])) && tp <= max_t) {
                if (model == 8) y += (v * dt) + (dt_alpha * random_stable(&R, alpha));
                else if (model == 7) y += ((v - (g * y)) * dt) + sqrt_st * random_gaussian(&R);
                else if (tdrift) y += (dr[ix] * dt) + (sqrt_st * random_gaussian(&R));
                else y += drift + (sqrt_st * random_gaussian(&R));
                tp += dt;
                ix++;
            }
            const size_t o = (size_t)n * n_sets + k;
            rts[o] = tp + t_nd;
            choices[o] = (model == 13) ? (-1) * sgn(y) : (zero_a ? (y < 0 ? 0 : 1) : sgn(y));
        }
    }
    free(b);
    free(dr);
    return 0;
}

/* race_model (:1253-1374) and lca (:1380-1518).  theta canonical layout per set (2*np + 5 floats):
 * v[np], a, z[np], g, b_inh, t, angle.  Bound: a (constant) or a + t_s * (-sin/cos(angle)) on the float32 time grid. */
int orc_sim_race_run(int lca, int np, int angle, const float* theta, int n_sets, int64_t n_samples, float dt, float max_t,
                     float s, uint64_t seed, float* rts, int32_t* choices)
{
    if (np < 2 || np > 4) return 1;
    rng_t R;
    rng_seed(&R, seed);
    const int n_par = 2 * np + 5;
    const float sqrt_st = (float)sqrt(dt) * s;
    const int nb = (int)((max_t / dt) + 1) + 2;
    float* b = (float*)malloc(sizeof(float) * (size_t)(nb + 1));
    if (!b) return 2;
    for (int k = 0; k < n_sets; k++) {
        const float* th = theta + (size_t)k * n_par;
        const float* v = th;
        const float a = th[np];
        const float* z = th + np + 1;
        const float g = th[2 * np + 1], binh = th[2 * np + 2], t = th[2 * np + 3], ang = th[2 * np + 4];
        for (int ix = 0; ix <= nb; ix++) {
            const float ts = (float)((double)ix * (double)dt);
            b[ix] = angle ? a + ts * (-sinf(ang) / cosf(ang)) : a;
        }
        for (int64_t n = 0; n < n_samples; n++) {
            float p[4], red[4];
            for (int j = 0; j < np; j++) p[j] = z[j] * b[0];
            float tp = 0.0f;
            int ix = 0;
            for (;;) {
                int fin = 0;
                for (int j = 0; j < np; j++) if (p[j] > b[ix]) fin = 1;
                if (fin || !(tp <= max_t)) break;
                if (lca) {
                    float sum = 0;
                    for (int j = 0; j < np; j++) sum += p[j];
                    for (int j = 0; j < np; j++) {
                        red[j] = (-1) * p[j] + sum;
                        p[j] += ((v[j] - (g * p[j]) - (binh * red[j])) * dt) + (sqrt_st * random_gaussian(&R));
                        p[j] = fmaxf(0.0f, p[j]);
                    }
                } else {
                    for (int j = 0; j < np; j++) p[j] += (v[j] * dt) + sqrt_st * random_gaussian(&R);
                }
                tp += dt;
                ix++;
            }
            int arg = 0;
            for (int j = 1; j < np; j++) if (p[j] > p[arg]) arg = j;
            const size_t o = (size_t)n * n_sets + k;
            rts[o] = tp + t;
            choices[o] = arg;
        }
    }
    free(b);
    return 0;
}

/* ddm_flexbound_par2 (:1667-1791).  theta per set (10 floats): v_h, v_l1, v_l2, a, z_h, z_l1, z_l2, t, b0, b1;
 * bound_kind 0 constant, 1 angle (b0), 2 Weibull (b0 = alpha, b1 = beta).  choice = (y_h < 0 ? 0 : 2) + (sign(y_l) >= 0). */
int orc_sim_par2_run(int bound_kind, const float* theta, int n_sets, int64_t n_samples, float dt, float max_t, float s,
                     uint64_t seed, float* rts, int32_t* choices)
{
    rng_t R;
    rng_seed(&R, seed);
    const float sqrt_st = (float)sqrt(dt) * s;
    const int nb = (int)((max_t / dt) + 1) + 2;
    float* b = (float*)malloc(sizeof(float) * (size_t)(nb + 1));
    if (!b) return 2;
    for (int k = 0; k < n_sets; k++) {
        const float* th = theta + (size_t)k * 10;
        const float a = th[3], t = th[7];
        for (int ix = 0; ix <= nb; ix++) {
            const float ts = (float)((double)ix * (double)dt);
            if (bound_kind == 1) b[ix] = a + ts * (-sinf(th[8]) / cosf(th[8]));
            else if (bound_kind == 2) b[ix] = (float)(a * exp(-pow((double)ts / th[9], (double)th[8])));
            else b[ix] = a;
        }
        for (int64_t n = 0; n < n_samples; n++) {
            float t_h = 0.0f, t_l = 0.0f;
            int ix = 0;
            float y_h = (-1) * b[0] + (th[4] * 2 * b[0]);
            while (y_h >= (-1) * b[ix] && y_h <= b[ix] && t_h <= max_t) {
                y_h += (th[0] * dt) + (sqrt_st * random_gaussian(&R));
                t_h += dt;
                ix++;
            }
            int choice;
            float y_l, v_l;
            if (sgn(y_h) < 0) { choice = 0; y_l = (-1) * b[0] + (th[5] * 2 * b[0]); v_l = th[1]; }
            else { choice = 2; y_l = (-1) * b[0] + (th[6] * 2 * b[0]); v_l = th[2]; }
            ix = 0;
            while (y_l >= (-1) * b[ix] && y_l <= b[ix] && t_l <= max_t) {
                y_l += (v_l * dt) + (sqrt_st * random_gaussian(&R));
                t_l += dt;
                ix++;
            }
            const size_t o = (size_t)n * n_sets + k;
            rts[o] = fmaxf(t_h, t_l) + t;
            if (sgn(y_l) >= 0) choice += 1;
            choices[o] = choice;
        }
    }
    free(b);
    return 0;
}

/* ddm_flexbound_seq2 (:1523-1661), ddm_flexbound_mic2_adj (:1797-1960), ddm_flexbound_tradeoff (:1966-2137).
 * kind 1 / 2 / 3.  theta per set (13 floats): v_h, v_l1, v_l2, a, z_h, z_l1, z_l2, t, b0, b1, d, b2, b3.
 * bound_kind 0 constant, 1 angle (b0), 2 Weibull (b0 = alpha, b1 = beta), 3 conflict-gamma (graveyard/boundary_functions.py:
 * 66-97: max(a + scale * gamma_pdf(t; shape, scale_gamma) + t * (-tan(theta)), 0) with b0 = theta, b1 = scale, b2 = shape,
 * b3 = scale_gamma).
 * Literal about the state the reference carries from one particle into the next: seq2 leaves y_l / v_l untouched on
 * its time-out and collapsed-bound branches, and the other two never rewrite bias_trace[n_h] (n_h = number of
 * high-dimension steps of the current particle).  Before the first particle that state is taken as zero. */
int orc_sim_hier_run(int kind, int bound_kind, const float* theta, int n_sets, int64_t n_samples, float dt, float max_t,
                     float s, uint64_t seed, float* rts, int32_t* choices)
{
    rng_t R;
    rng_seed(&R, seed);
    const float sqrt_st = (float)sqrt(dt) * s;
    const int num_draws = (int)((max_t / dt) + 1);
    const int nb = num_draws + 2;
    float* b = (float*)malloc(sizeof(float) * (size_t)(nb + 1));
    float* trace = (float*)calloc((size_t)(nb + 1), sizeof(float));
    if (!b || !trace) { free(b); free(trace); return 2; }
    float y_l = 0.0f, v_l = 0.0f;
    for (int k = 0; k < n_sets; k++) {
        const float* th = theta + (size_t)k * 13;
        const float a = th[3], t = th[7], d = th[10];
        for (int ix = 0; ix <= nb; ix++) {
            const float ts = (float)((double)ix * (double)dt);
            if (bound_kind == 1) b[ix] = a + ts * (-sinf(th[8]) / cosf(th[8]));
            else if (bound_kind == 2) b[ix] = (float)(a * exp(-pow((double)ts / th[9], (double)th[8])));
            else if (bound_kind == 3) {
                const double shape = th[11], sg = th[12], x = (double)ts / sg;
                double pdf;
                if (x > 0) pdf = exp((shape - 1.0) * log(x) - x - lgamma(shape)) / sg;
                else pdf = (shape > 1.0) ? 0.0 : (shape == 1.0 ? 1.0 / sg : INFINITY);
                const double v = (double)a + (double)th[9] * pdf + (double)ts * (-sin((double)th[8]) / cos((double)th[8]));
                b[ix] = (float)(v > 0 ? v : 0);
            } else b[ix] = a;
        }
        for (int64_t n = 0; n < n_samples; n++) {
            const size_t o = (size_t)n * n_sets + k;
            int choice = 0, ix = 0;
            float y_h = (-1) * b[0] + (th[4] * 2 * b[0]);
            if (kind == 1) {
                float tp = 0.0f;
                while (y_h >= (-1) * b[ix] && y_h <= b[ix] && tp <= max_t) {
                    y_h += (th[0] * dt) + (sqrt_st * random_gaussian(&R));
                    tp += dt;
                    ix++;
                }
                if (tp >= max_t) {
                    if (random_uniform(&R) > 0.5f) choice += 1;
                } else {
                    const int lower = sgn(y_h) < 0;
                    const float z_l = lower ? th[5] : th[6];
                    choice = lower ? 0 : 2;
                    if (b[ix] <= 0) {
                        if (random_uniform(&R) < z_l) choice += 1;
                    } else {
                        y_l = (-1) * b[ix] + (z_l * 2 * b[ix]);
                        v_l = lower ? th[1] : th[2];
                    }
                }
                while (y_l >= (-1) * b[ix] && y_l <= b[ix] && tp <= max_t) {
                    y_l += (v_l * dt) + (sqrt_st * random_gaussian(&R));
                    tp += dt;
                    ix++;
                }
                rts[o] = tp + t;
                if (sgn(y_l) >= 0) choice += 1;
                choices[o] = choice;
                continue;
            }
            float t_h = 0.0f, t_l = 0.0f;
            trace[0] = (y_h + b[0]) / (2 * b[0]);
            while (y_h >= (-1) * b[ix] && y_h <= b[ix] && t_h <= max_t) {
                y_h += (th[0] * dt) + (sqrt_st * random_gaussian(&R));
                trace[ix] = (y_h + b[ix]) / (2 * b[ix]);
                t_h += dt;
                ix++;
            }
            if (random_uniform(&R) <= ((y_h + b[ix]) / (2 * b[ix]))) choice += 2;
            if (choice == 2) {
                y_l = (-1) * b[0] + (th[6] * 2 * b[0]);
                v_l = th[2];
                for (int j = ix + 1; j < num_draws; j++) trace[j] = 1.0f;
            } else {
                y_l = (-1) * b[0] + (th[5] * 2 * b[0]);
                v_l = th[1];
                for (int j = ix + 1; j < num_draws; j++) trace[j] = 0.0f;
                for (int j = 0; j < num_draws; j++) trace[j] = 1.0f - trace[j];
            }
            ix = 0;
            while (y_l >= (-1) * b[ix] && y_l <= b[ix] && t_l <= max_t) {
                const int before = (trace[ix] < 1) && (trace[ix] > 0);
                if (kind == 2) {
                    if (before) y_l += ((v_l * trace[ix] * d) * dt);
                    else y_l += (v_l * dt);
                    y_l += (sqrt_st * random_gaussian(&R));
                } else {
                    const float pos = (float)((1 + ((double)d * ((double)trace[ix] - 1.0))) / (double)(2 - d));
                    const float g = random_gaussian(&R);
                    if (before) {
                        y_l += pos * (v_l * dt);
                        y_l += pos * sqrt_st * g;
                    } else {
                        y_l += (v_l * dt);
                        y_l += sqrt_st * g;
                    }
                }
                t_l += dt;
                ix++;
            }
            rts[o] = fmaxf(t_h, t_l) + t;
            if (random_uniform(&R) <= ((y_l + b[ix]) / (2 * b[ix]))) choice += 1;
            choices[o] = choice;
        }
    }
    free(b);
    free(trace);
    return 0;
}
