/*
 * oracle/wfpt_oracle.c -- CPU restatement of HDDM's WFPT likelihood path.  TEST INFRASTRUCTURE ONLY.
 *
 * Nothing under hddm_b200/ may link, load or call this file: it is the checker that tests/, bench.py's
 * cpu_baseline leg and __graft_entry__.smoke() compare the CUDA path against.  Plain C, FP64, no contraction
 * (compile with -O2 -ffp-contract=off), one thread.
 *
 * Each function names the reference lines it restates (paths relative to /root/reference):
 *   src/pdf.pxi        ftt_01w :28-65, prob_ub :67-72, pdf :74-85, pdf_sv :87-102, full_pdf :104-146
 *   src/integrate.pxi  simpson_1D :12-45, simpson_2D :47-70, adaptiveSimpsonsAux :72-112,
 *                      adaptiveSimpsons_1D :114-141, adaptiveSimpsonsAux_2D :143-178, adaptiveSimpsons_2D :181-206
 *   src/wfpt.pyx       pdf_array :32-48, wiener_like :54-76, wiener_like_rlddm :78-153, wiener_like_rl :270-354,
 *                      wiener_like_multi :357-387, wiener_like_multi_rlddm :390-433, gen_rts_from_cdf :517-548 (cdf part),
 *                      wiener_like_contaminant :551-580
 *
 * Parity pinning: tests/test_oracle_golden.py checks this file against the reference's own golden vectors
 * (hddm/tests/matlab_values.py, test_likelihoods.py known answers) and, bit for bit, against the reference's
 * Cython compiled unmodified into oracle/_ref (fixtures in tests/golden/ are generated from that build).
 *
 * Notes on how Cython lowers the reference's arithmetic (verified in the generated C++ of the oracle/_ref build):
 *   x**2 / pow(x,2)  -> pow(x, 2.0), which gcc folds to x*x;   M_PI**2 -> pow(M_PI, 2.0) == fl(M_PI*M_PI);
 *   pow(tt,3) stays a libm call;  `cdivision=True` gives C division;  no -march flags -> no FMA contraction.
 */
#include <math.h>
#include <stddef.h>
#include <stdint.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

typedef struct {
    double err;        /* series truncation error bound (pdf.pxi `err`) */
    int n_st;          /* fixed grid: #intervals over t;  adaptive: max recursion depth over t */
    int n_sz;          /* same for z */
    int use_adaptive;  /* pdf.pxi:106 */
    double simps_err;  /* adaptive Simpson tolerance */
} orc_opts;

/* optional instrumentation: number of pdf_sv node evaluations since last reset (survey 8(d): nodes/trial) */
static long long g_nodes = 0;
long long orc_nodes_get(void) { return g_nodes; }
void orc_nodes_reset(void) { g_nodes = 0; }

/* ---- pdf.pxi:28-65 ------------------------------------------------------------------------------------ */
/* f(tt | 0, 1, w): Navarro & Fuss (2009).  The number of series terms and the choice between the small-time and
 * the large-time expansion are fixed from (tt, err) alone. `Kout`, `small_out` report that decision (may be NULL). */
double orc_ftt_01w_info(double tt, double w, double err, int *Kout, int *small_out)
{
    double kl, ks, p = 0.0;
    int k, K;

    /* large-time term count (:36-40) */
    if (M_PI * tt * err < 1) {
        kl = sqrt(-2 * log(M_PI * tt * err) / ((M_PI * M_PI) * tt));
        double floor_l = 1. / (M_PI * sqrt(tt));
        kl = kl < floor_l ? floor_l : kl; /* std::max(a,b) = (a<b)?b:a */
    } else {
        kl = 1. / (M_PI * sqrt(tt));
    }
    /* small-time term count (:43-47) */
    if (2 * sqrt(2 * M_PI * tt) * err < 1) {
        ks = 2 + sqrt(-2 * tt * log(2 * sqrt(2 * M_PI * tt) * err));
        double floor_s = sqrt(tt) + 1;
        ks = ks < floor_s ? floor_s : ks;
    } else {
        ks = 2;
    }

    if (ks < kl) { /* small-time expansion (:51-57) */
        K = (int)ceil(ks);
        int lower = (int)(-floor((K - 1) / 2.));
        int upper = (int)(ceil((K - 1) / 2.));
        for (k = lower; k <= upper; k++) {
            double u = w + 2 * k;
            p += u * exp(-(u * u) / 2 / tt);
        }
        p /= sqrt(2 * M_PI * pow(tt, 3));
        if (small_out) *small_out = 1;
    } else { /* large-time expansion (:59-63) */
        K = (int)ceil(kl);
        for (k = 1; k <= K; k++) {
            double kk = (double)k;
            p += k * exp(-(kk * kk) * (M_PI * M_PI) * tt / 2) * sin(k * M_PI * w);
        }
        p *= M_PI;
        if (small_out) *small_out = 0;
    }
    if (Kout) *Kout = K;
    return p;
}

double orc_ftt_01w(double tt, double w, double err) { return orc_ftt_01w_info(tt, w, err, NULL, NULL); }

/* ---- pdf.pxi:67-72 ------------------------------------------------------------------------------------ */
double orc_prob_ub(double v, double a, double z)
{
    if (v == 0) return z;
    return (exp(-2 * a * z * v) - 1) / (exp(-2 * a * v) - 1);
}

/* ---- pdf.pxi:74-85 ------------------------------------------------------------------------------------ */
double orc_pdf(double x, double v, double a, double w, double err)
{
    if (x <= 0) return 0;
    double tt = x / (a * a);
    double p = orc_ftt_01w(tt, w, err);
    return p * exp(-v * a * w - (v * v) * x / 2.) / (a * a);
}

/* ---- pdf.pxi:87-102 ----------------------------------------------------------------------------------- */
double orc_pdf_sv(double x, double v, double sv, double a, double z, double err)
{
    g_nodes++;
    if (x <= 0) return 0;
    if (sv == 0) return orc_pdf(x, v, a, z, err);
    double tt = x / (a * a);
    double p = orc_ftt_01w(tt, z, err);
    double azsv = a * z * sv;
    return exp(log(p) + (azsv * azsv - 2 * a * v * z - (v * v) * x) / (2 * (sv * sv) * x + 2)) / sqrt((sv * sv) * x + 1) /
           (a * a);
}

/* ---- integrate.pxi:12-45 ------------------------------------------------------------------------------ */
static double simpson_1D(double x, double v, double sv, double a, double z, double t, double err, double lb_z,
                         double ub_z, int n_sz, double lb_t, double ub_t, int n_st)
{
    double ht, hz;
    int n = n_st > n_sz ? n_st : n_sz;
    if (n_st == 0) { /* integrate over z, t fixed */
        hz = (ub_z - lb_z) / n;
        ht = 0;
        lb_t = t;
        ub_t = t;
    } else { /* integrate over t, z fixed */
        hz = 0;
        ht = (ub_t - lb_t) / n;
        lb_z = z;
        ub_z = z;
    }
    double S = orc_pdf_sv(x - lb_t, v, sv, a, lb_z, err);
    double y = 0; /* reference leaves y uninitialised when n==0 (Appendix A.12); never reached with sz|st>0 */
    for (int i = 1; i <= n; i++) {
        double z_tag = lb_z + hz * i;
        double t_tag = lb_t + ht * i;
        y = orc_pdf_sv(x - t_tag, v, sv, a, z_tag, err);
        if (i & 1)
            S += (4 * y);
        else
            S += (2 * y);
    }
    S = S - y;
    S = S / ((ub_t - lb_t) + (ub_z - lb_z));
    return ((ht + hz) * S / 3);
}

/* ---- integrate.pxi:47-70 ------------------------------------------------------------------------------ */
static double simpson_2D(double x, double v, double sv, double a, double z, double t, double err, double lb_z,
                         double ub_z, int n_sz, double lb_t, double ub_t, int n_st)
{
    (void)t;
    double ht = (ub_t - lb_t) / n_st;
    double S = simpson_1D(x, v, sv, a, z, lb_t, err, lb_z, ub_z, n_sz, 0, 0, 0);
    double y = 0;
    for (int i_t = 1; i_t <= n_st; i_t++) {
        double t_tag = lb_t + ht * i_t;
        y = simpson_1D(x, v, sv, a, z, t_tag, err, lb_z, ub_z, n_sz, 0, 0, 0);
        if (i_t & 1)
            S += (4 * y);
        else
            S += (2 * y);
    }
    S = S - y;
    S = S / (ub_t - lb_t);
    return (ht * S / 3);
}

/* ---- integrate.pxi:72-112 ----------------------------------------------------------------------------- */
static double adaptive_aux(double x, double v, double sv, double a, double z, double t, double pdf_err, double lb_z,
                           double ub_z, double lb_t, double ub_t, double ZT, double simps_err, double S, double f_beg,
                           double f_end, double f_mid, int bottom)
{
    double z_c, z_d, z_e, t_c, t_d, t_e, h;
    if ((ub_t - lb_t) == 0) { /* over z */
        h = ub_z - lb_z;
        z_c = (ub_z + lb_z) / 2.;
        z_d = (lb_z + z_c) / 2.;
        z_e = (z_c + ub_z) / 2.;
        t_c = t;
        t_d = t;
        t_e = t;
    } else { /* over t */
        h = ub_t - lb_t;
        t_c = (ub_t + lb_t) / 2.;
        t_d = (lb_t + t_c) / 2.;
        t_e = (t_c + ub_t) / 2.;
        z_c = z;
        z_d = z;
        z_e = z;
    }
    double fd = orc_pdf_sv(x - t_d, v, sv, a, z_d, pdf_err) / ZT;
    double fe = orc_pdf_sv(x - t_e, v, sv, a, z_e, pdf_err) / ZT;
    double Sleft = (h / 12) * (f_beg + 4 * fd + f_mid);
    double Sright = (h / 12) * (f_mid + 4 * fe + f_end);
    double S2 = Sleft + Sright;
    if (bottom <= 0 || fabs(S2 - S) <= 15 * simps_err) return S2 + (S2 - S) / 15;
    double left = adaptive_aux(x, v, sv, a, z, t, pdf_err, lb_z, z_c, lb_t, t_c, ZT, simps_err / 2, Sleft, f_beg, f_mid,
                               fd, bottom - 1);
    double right = adaptive_aux(x, v, sv, a, z, t, pdf_err, z_c, ub_z, t_c, ub_t, ZT, simps_err / 2, Sright, f_mid,
                                f_end, fe, bottom - 1);
    return left + right;
}

/* ---- integrate.pxi:114-141 ---------------------------------------------------------------------------- */
static double adaptive_1D(double x, double v, double sv, double a, double z, double t, double pdf_err, double lb_z,
                          double ub_z, double lb_t, double ub_t, double simps_err, int max_depth)
{
    double h;
    if ((ub_t - lb_t) == 0) {
        lb_t = t;
        ub_t = t;
        h = ub_z - lb_z;
    } else {
        h = (ub_t - lb_t);
        lb_z = z;
        ub_z = z;
    }
    double ZT = h;
    double c_t = (lb_t + ub_t) / 2.;
    double c_z = (lb_z + ub_z) / 2.;
    double f_beg = orc_pdf_sv(x - lb_t, v, sv, a, lb_z, pdf_err) / ZT;
    double f_end = orc_pdf_sv(x - ub_t, v, sv, a, ub_z, pdf_err) / ZT;
    double f_mid = orc_pdf_sv(x - c_t, v, sv, a, c_z, pdf_err) / ZT;
    double S = (h / 6) * (f_beg + 4 * f_mid + f_end);
    return adaptive_aux(x, v, sv, a, z, t, pdf_err, lb_z, ub_z, lb_t, ub_t, ZT, simps_err, S, f_beg, f_end, f_mid,
                        max_depth);
}

/* ---- integrate.pxi:143-178 ---------------------------------------------------------------------------- */
static double adaptive_aux_2D(double x, double v, double sv, double a, double z, double t, double pdf_err,
                              double err_1d, double lb_z, double ub_z, double lb_t, double ub_t, double st,
                              double err_2d, double S, double f_beg, double f_end, double f_mid, int depth_sz,
                              int bottom)
{
    double t_c = (ub_t + lb_t) / 2.;
    double t_d = (lb_t + t_c) / 2.;
    double t_e = (t_c + ub_t) / 2.;
    double h = ub_t - lb_t;
    double fd = adaptive_1D(x, v, sv, a, z, t_d, pdf_err, lb_z, ub_z, 0, 0, err_1d, depth_sz) / st;
    double fe = adaptive_1D(x, v, sv, a, z, t_e, pdf_err, lb_z, ub_z, 0, 0, err_1d, depth_sz) / st;
    double Sleft = (h / 12) * (f_beg + 4 * fd + f_mid);
    double Sright = (h / 12) * (f_mid + 4 * fe + f_end);
    double S2 = Sleft + Sright;
    if (bottom <= 0 || fabs(S2 - S) <= 15 * err_2d) return S2 + (S2 - S) / 15;
    double left = adaptive_aux_2D(x, v, sv, a, z, t, pdf_err, err_1d, lb_z, ub_z, lb_t, t_c, st, err_2d / 2, Sleft,
                                  f_beg, f_mid, fd, depth_sz, bottom - 1);
    double right = adaptive_aux_2D(x, v, sv, a, z, t, pdf_err, err_1d, lb_z, ub_z, t_c, ub_t, st, err_2d / 2, Sright,
                                   f_mid, f_end, fe, depth_sz, bottom - 1);
    return left + right;
}

/* ---- integrate.pxi:181-206 ---------------------------------------------------------------------------- */
static double adaptive_2D(double x, double v, double sv, double a, double z, double t, double pdf_err, double lb_z,
                          double ub_z, double lb_t, double ub_t, double simps_err, int depth_sz, int depth_st)
{
    double h = (ub_t - lb_t);
    double st = (ub_t - lb_t);
    double err_1d = simps_err, err_2d = simps_err;
    double f_beg = adaptive_1D(x, v, sv, a, z, lb_t, pdf_err, lb_z, ub_z, 0, 0, err_1d, depth_sz) / st;
    double f_end = adaptive_1D(x, v, sv, a, z, ub_t, pdf_err, lb_z, ub_z, 0, 0, err_1d, depth_sz) / st;
    double f_mid = adaptive_1D(x, v, sv, a, z, (lb_t + ub_t) / 2, pdf_err, lb_z, ub_z, 0, 0, err_1d, depth_sz) / st;
    double S = (h / 6) * (f_beg + 4 * f_mid + f_end);
    return adaptive_aux_2D(x, v, sv, a, z, t, pdf_err, err_1d, lb_z, ub_z, lb_t, ub_t, st, err_2d, S, f_beg, f_end,
                           f_mid, depth_sz, depth_st);
}

/* ---- pdf.pxi:104-146 ---------------------------------------------------------------------------------- */
double orc_full_pdf(double x, double v, double sv, double a, double z, double sz, double t, double st, double err,
                    int n_st, int n_sz, int use_adaptive, double simps_err)
{
    if ((z < 0) || (z > 1) || (a < 0) || (t < 0) || (st < 0) || (sv < 0) || (sz < 0) || (sz > 1) ||
        ((fabs(x) - (t - st / 2.)) < 0) || (z + sz / 2. > 1) || (z - sz / 2. < 0) || (t - st / 2. < 0))
        return 0;

    if (x > 0) {
        v = -v;
        z = 1. - z;
    }
    x = fabs(x);
    if (st < 1e-3) st = 0;
    if (sz < 1e-3) sz = 0;

    if (sz == 0) {
        if (st == 0) return orc_pdf_sv(x - t, v, sv, a, z, err);
        if (use_adaptive > 0)
            return adaptive_1D(x, v, sv, a, z, t, err, z, z, t - st / 2., t + st / 2., simps_err, n_st);
        return simpson_1D(x, v, sv, a, z, t, err, z, z, 0, t - st / 2., t + st / 2., n_st);
    }
    if (st == 0) {
        if (use_adaptive)
            return adaptive_1D(x, v, sv, a, z, t, err, z - sz / 2., z + sz / 2., t, t, simps_err, n_sz);
        return simpson_1D(x, v, sv, a, z, t, err, z - sz / 2., z + sz / 2., n_sz, t, t, 0);
    }
    if (use_adaptive)
        return adaptive_2D(x, v, sv, a, z, t, err, z - sz / 2., z + sz / 2., t - st / 2., t + st / 2., simps_err, n_sz,
                           n_st);
    return simpson_2D(x, v, sv, a, z, t, err, z - sz / 2., z + sz / 2., n_sz, t - st / 2., t + st / 2., n_st);
}

/* ---- wfpt.pyx:32-48 ----------------------------------------------------------------------------------- */
/* x may be strided (stride in elements), as the Cython buffer accepts non-contiguous arrays. */
void orc_pdf_array(const double *x, ptrdiff_t n, ptrdiff_t stride, double v, double sv, double a, double z, double sz,
                   double t, double st, double err, int logp, int n_st, int n_sz, int use_adaptive, double simps_err,
                   double p_outlier, double w_outlier, double *y)
{
    double wp = w_outlier * p_outlier;
    for (ptrdiff_t i = 0; i < n; i++) {
        double p = orc_full_pdf(x[i * stride], v, sv, a, z, sz, t, st, err, n_st, n_sz, use_adaptive, simps_err);
        p = p * (1 - p_outlier) + wp;
        y[i] = logp ? log(p) : p;
    }
}

/* ---- wfpt.pyx:54-76 ----------------------------------------------------------------------------------- */
double orc_wiener_like(const double *x, ptrdiff_t n, ptrdiff_t stride, double v, double sv, double a, double z,
                       double sz, double t, double st, double err, int n_st, int n_sz, int use_adaptive,
                       double simps_err, double p_outlier, double w_outlier)
{
    double sum_logp = 0;
    double wp_outlier = w_outlier * p_outlier;
    if (!((p_outlier >= 0) & (p_outlier <= 1))) return -INFINITY;
    for (ptrdiff_t i = 0; i < n; i++) {
        double p = orc_full_pdf(x[i * stride], v, sv, a, z, sz, t, st, err, n_st, n_sz, use_adaptive, simps_err);
        p = p * (1 - p_outlier) + wp_outlier;
        if (p == 0) return -INFINITY;
        sum_logp += log(p);
    }
    return sum_logp;
}

/* Batched form used as the CPU arm of the benchmark: params[P][8] = v,sv,a,z,sz,t,st,p_outlier (row-major).
 * Exactly P independent orc_wiener_like calls -- this is how a caller of the reference would evaluate a grid. */
void orc_wiener_like_batch(const double *x, ptrdiff_t n, const double *params, ptrdiff_t P, double err, int n_st,
                           int n_sz, int use_adaptive, double simps_err, double w_outlier, double *out)
{
    for (ptrdiff_t j = 0; j < P; j++) {
        const double *q = params + 8 * j;
        out[j] = orc_wiener_like(x, n, 1, q[0], q[1], q[2], q[3], q[4], q[5], q[6], err, n_st, n_sz, use_adaptive,
                                 simps_err, q[7], w_outlier);
    }
}

/* ---- wfpt.pyx:357-387 --------------------------------------------------------------------------------- */
/* per_trial[k] (k = v,sv,a,z,sz,t,st) is either NULL (use scalars[k]) or an array of length n. */
double orc_wiener_like_multi(const double *x, ptrdiff_t n, const double *const per_trial[7], const double scalars[7],
                             double err, int n_st, int n_sz, int use_adaptive, double simps_err, double p_outlier,
                             double w_outlier)
{
    double sum_logp = 0, p;
    double wp_outlier = w_outlier * p_outlier;
    for (ptrdiff_t i = 0; i < n; i++) {
        double q[7];
        for (int k = 0; k < 7; k++) q[k] = per_trial[k] ? per_trial[k][i] : scalars[k];
        if (fabs(x[i]) != 999.) {
            p = orc_full_pdf(x[i], q[0], q[1], q[2], q[3], q[4], q[5], q[6], err, n_st, n_sz, use_adaptive, simps_err);
            p = p * (1 - p_outlier) + wp_outlier;
        } else if (x[i] == 999.) {
            p = orc_prob_ub(q[0], q[2], q[3]);
        } else {
            p = 1 - orc_prob_ub(q[0], q[2], q[3]);
        }
        sum_logp += log(p);
    }
    return sum_logp;
}

/* ---- wfpt.pyx:78-153 ---------------------------------------------------------------------------------- */
/* RL learning-rate transform with the reference's truncated base (Appendix A.8). */
static double rl_rate(double alpha) { return pow(2.718281828459, alpha) / (1 + pow(2.718281828459, alpha)); }

/* `order` lists trial indices grouped by ascending split_by value, original order inside a group (what
 * np.unique + boolean masks produce); group g covers order[starts[g] .. starts[g+1]). */
double orc_wiener_like_rlddm(const double *x, const int64_t *response, const double *feedback, const int64_t *order,
                             const int64_t *starts, ptrdiff_t n_groups, double q, double alpha, double pos_alpha,
                             double v, double sv, double a, double z, double sz, double t, double st, double err,
                             int n_st, int n_sz, int use_adaptive, double simps_err, double p_outlier, double w_outlier)
{
    double sum_logp = 0;
    double wp_outlier = w_outlier * p_outlier;
    double pos_alfa, alfa, qs[2];
    if (!((p_outlier >= 0) & (p_outlier <= 1))) return -INFINITY;
    pos_alfa = (pos_alpha == 100.00) ? alpha : pos_alpha;
    for (ptrdiff_t g = 0; g < n_groups; g++) {
        const int64_t *idx = order + starts[g];
        ptrdiff_t s_size = (ptrdiff_t)(starts[g + 1] - starts[g]);
        qs[0] = q;
        qs[1] = q;
        /* first trial: update only */
        alfa = (feedback[idx[0]] > qs[response[idx[0]]]) ? rl_rate(pos_alfa) : rl_rate(alpha);
        qs[response[idx[0]]] = qs[response[idx[0]]] + alfa * (feedback[idx[0]] - qs[response[idx[0]]]);
        for (ptrdiff_t i = 1; i < s_size; i++) {
            int64_t j = idx[i];
            double p = orc_full_pdf(x[j], ((qs[1] - qs[0]) * v), sv, a, z, sz, t, st, err, n_st, n_sz, use_adaptive,
                                    simps_err);
            p = p * (1 - p_outlier) + wp_outlier;
            if (p == 0) return -INFINITY;
            sum_logp += log(p);
            alfa = (feedback[j] > qs[response[j]]) ? rl_rate(pos_alfa) : rl_rate(alpha);
            qs[response[j]] = qs[response[j]] + alfa * (feedback[j] - qs[response[j]]);
        }
    }
    return sum_logp;
}

/* ---- wfpt.pyx:270-354 --------------------------------------------------------------------------------- */
double orc_wiener_like_rl(const int64_t *response, const double *feedback, const int64_t *order, const int64_t *starts,
                          ptrdiff_t n_groups, double q, double alpha, double pos_alpha, double v, double z,
                          double p_outlier, double w_outlier)
{
    double sum_logp = 0;
    double wp_outlier = w_outlier * p_outlier;
    double pos_alfa, alfa, qs[2], p;
    if (!((p_outlier >= 0) & (p_outlier <= 1))) return -INFINITY;
    pos_alfa = (pos_alpha == 100.00) ? alpha : pos_alpha;
    for (ptrdiff_t g = 0; g < n_groups; g++) {
        const int64_t *idx = order + starts[g];
        ptrdiff_t s_size = (ptrdiff_t)(starts[g + 1] - starts[g]);
        qs[0] = q;
        qs[1] = q;
        alfa = (feedback[idx[0]] > qs[response[idx[0]]]) ? rl_rate(pos_alfa) : rl_rate(alpha);
        qs[response[idx[0]]] = qs[response[idx[0]]] + alfa * (feedback[idx[0]] - qs[response[idx[0]]]);
        for (ptrdiff_t i = 1; i < s_size; i++) {
            int64_t j = idx[i];
            double drift = (qs[1] - qs[0]) * v;
            if (drift == 0) {
                p = 0.5;
            } else if (response[j] == 1) {
                p = (pow(2.718281828459, -2 * z * drift) - 1) / (pow(2.718281828459, -2 * drift) - 1);
            } else {
                p = 1 - (pow(2.718281828459, -2 * z * drift) - 1) / (pow(2.718281828459, -2 * drift) - 1);
            }
            p = p * (1 - p_outlier) + wp_outlier;
            if (p == 0) return -INFINITY;
            sum_logp += log(p);
            alfa = (feedback[j] > qs[response[j]]) ? rl_rate(pos_alfa) : rl_rate(alpha);
            qs[response[j]] = qs[response[j]] + alfa * (feedback[j] - qs[response[j]]);
        }
    }
    return sum_logp;
}

/* ---- wfpt.pyx:390-433 --------------------------------------------------------------------------------- */
/* per_trial[0..6] = v,sv,a,z,sz,t,st ; per_trial[7] = alpha.  Trials run in array order; Q resets when
 * split_by changes between consecutive trials; no first-trial skip and no early exit (unlike wiener_like_rlddm). */
double orc_wiener_like_multi_rlddm(const double *x, const int64_t *response, const double *feedback,
                                   const int64_t *split_by, ptrdiff_t n, double q, const double *const per_trial[8],
                                   const double scalars[8], double err, int n_st, int n_sz, int use_adaptive,
                                   double simps_err, double p_outlier, double w_outlier)
{
    double sum_logp = 0;
    double wp_outlier = w_outlier * p_outlier;
    double qs[2] = {q, q};
    for (ptrdiff_t i = 0; i < n; i++) {
        double c[8];
        for (int k = 0; k < 8; k++) c[k] = per_trial[k] ? per_trial[k][i] : scalars[k];
        if (i != 0 && split_by[i] != split_by[i - 1]) {
            qs[0] = q;
            qs[1] = q;
        }
        double p = orc_full_pdf(x[i], c[0] * (qs[1] - qs[0]), c[1], c[2], c[3], c[4], c[5], c[6], err, n_st, n_sz,
                                use_adaptive, simps_err);
        p = p * (1 - p_outlier) + wp_outlier;
        sum_logp += log(p);
        double alfa = rl_rate(c[7]);
        qs[response[i]] = qs[response[i]] + alfa * (feedback[i] - qs[response[i]]);
    }
    return sum_logp;
}

/* ---- wfpt.pyx:551-580 --------------------------------------------------------------------------------- */
double orc_wiener_like_contaminant(const double *x, const int32_t *cont_x, ptrdiff_t n, double v, double sv, double a,
                                   double z, double sz, double t, double st, double t_min, double t_max, double err,
                                   int n_st, int n_sz, int use_adaptive, double simps_err)
{
    double sum_logp = 0;
    int n_cont = 0;
    for (ptrdiff_t i = 0; i < n; i++) n_cont += cont_x[i];
    for (ptrdiff_t i = 0; i < n; i++) {
        if (cont_x[i] == 0) {
            double p = orc_full_pdf(x[i], v, sv, a, z, sz, t, st, err, n_st, n_sz, use_adaptive, simps_err);
            if (p == 0) return -INFINITY;
            sum_logp += log(p);
        }
    }
    sum_logp += n_cont * log(0.5 * 1. / (t_max - t_min));
    return sum_logp;
}

/* ---- wfpt.pyx:517-531 (deterministic half of gen_rts_from_cdf) ------------------------------------------ */
/* l_cdf[0]=0; l_cdf[i] = l_cdf[i-1] + full_pdf(x[i], v,sv,a,z,sz, 0,0, 1e-4) with full_pdf's defaults; then /= last. */
void orc_cdf_grid(const double *x, ptrdiff_t n, double v, double sv, double a, double z, double sz, double *l_cdf)
{
    l_cdf[0] = 0;
    for (ptrdiff_t i = 1; i < n; i++) l_cdf[i] = l_cdf[i - 1] + orc_full_pdf(x[i], v, sv, a, z, sz, 0, 0, 1e-4, 2, 2, 1, 1e-3);
    double last = l_cdf[n - 1];
    for (ptrdiff_t i = 0; i < n; i++) l_cdf[i] /= last;
}
