// Operation-for-operation device evaluation of the WFPT density ("faithful" variant).
//
// This header must be compiled with --fmad=false: every product and sum below is rounded separately, in the
// order the reference writes it, so that the series term count K = ceil(ks|kl), the small/large-time branch,
// the Simpson node positions (lb + h*i) and the outlier mix p*(1-p_outlier)+w*p_outlier see the same doubles as
// the reference's x86-64 build (which has no FMA contraction).  The only residual differences are the <=1-2 ulp
// of CUDA's exp/log/sin versus glibc's.
//
// What it evaluates (reference lines, paths relative to /root/reference/src):
//   ftt_01w  pdf.pxi:28-65     pdf      pdf.pxi:74-85      pdf_sv  pdf.pxi:87-102    full_pdf pdf.pxi:104-146
//   fixed-grid Simpson integrate.pxi:12-70; adaptive Simpson integrate.pxi:72-206 (here: explicit stacks, no recursion)
#pragma once
#include "common.cuh"

namespace wfpt {
namespace faithful {

__device__ __forceinline__ double max_std(double a, double b) { return (a < b) ? b : a; }  // std::max

// f(tt | 0, 1, w).  K and the expansion are fixed by (tt, err) alone.
static __device__ __noinline__ double ftt_01w(double tt, double w, double err)
{
    double kl, ks, p = 0.0;
    if (kPi * tt * err < 1) {
        kl = sqrt(-2 * log(kPi * tt * err) / (kPiSq * tt));
        kl = max_std(kl, 1. / (kPi * sqrt(tt)));
    } else {
        kl = 1. / (kPi * sqrt(tt));
    }
    if (2 * sqrt(2 * kPi * tt) * err < 1) {
        ks = 2 + sqrt(-2 * tt * log(2 * sqrt(2 * kPi * tt) * err));
        ks = max_std(ks, sqrt(tt) + 1);
    } else {
        ks = 2;
    }
    if (ks < kl) {
        const int K = (int)ceil(ks);
        const int lower = (int)(-floor((K - 1) / 2.));
        const int upper = (int)(ceil((K - 1) / 2.));
        for (int k = lower; k <= upper; k++) {
            const double u = w + 2 * k;
            p += u * exp(-(u * u) / 2 / tt);
        }
        p /= sqrt(2 * kPi * (tt * tt * tt));
    } else {
        const int K = (int)ceil(kl);
        for (int k = 1; k <= K; k++) {
            const double kk = (double)k;
            p += k * exp(-(kk * kk) * kPiSq * tt / 2) * sin(k * kPi * w);
        }
        p *= kPi;
    }
    return p;
}

__device__ __forceinline__ double pdf(double x, double v, double a, double w, double err)
{
    if (x <= 0) return 0;
    const double tt = x / (a * a);
    const double p = ftt_01w(tt, w, err);
    return p * exp(-v * a * w - (v * v) * x / 2.) / (a * a);
}

// out of line: it is reached from several quadrature walkers, and one copy keeps the slow paths' code footprint small
static __device__ __noinline__ double pdf_sv(double x, double v, double sv, double a, double z, double err)
{
    if (x <= 0) return 0;
    if (sv == 0) return pdf(x, v, a, z, err);
    const double tt = x / (a * a);
    const double p = ftt_01w(tt, z, err);
    const double azsv = a * z * sv;
    return exp(log(p) + (azsv * azsv - 2 * a * v * z - (v * v) * x) / (2 * (sv * sv) * x + 2)) / sqrt((sv * sv) * x + 1) /
           (a * a);
}

// Everything a quadrature node needs besides its own (t, z) position.
struct Trial {
    double x, v, sv, a, err;
};

// Integrand along one axis.  over_z: s is a z position at fixed t;  else s is a t position at fixed z.
struct Axis {
    Trial q;
    double fixed;
    bool over_z;
    __device__ __forceinline__ double operator()(double s) const
    {
        return over_z ? pdf_sv(q.x - fixed, q.v, q.sv, q.a, s, q.err) : pdf_sv(q.x - s, q.v, q.sv, q.a, fixed, q.err);
    }
};

// integrate.pxi:12-45 with its 1-4-2-...-1 accumulation order.  n intervals over [lb, ub].
template <class F>
__device__ __forceinline__ double simpson_fixed(const F& f, double lb, double ub, int n)
{
    const double h = (ub - lb) / n;
    double S = 0;
    double y = 0;
    for (int i = 0; i <= n; i++) {   // node 0 is f(lb) with weight 1; one evaluation site for all nodes
        y = f(i == 0 ? lb : lb + h * i);
        S = (i == 0) ? y : S + ((i & 1) ? (4 * y) : (2 * y));
    }
    S = S - y;
    S = S / (ub - lb);
    return (h * S / 3);
}

// integrate.pxi:72-141 (and :143-206 when F is itself an integral): adaptive Simpson on f/ZT over [lb, ub], the
// recursion unrolled into an explicit post-order walk so that the additions left + right happen in the reference's
// order.  Frame d keeps the right child's arguments and, once known, the left child's value.
template <class F>
__device__ __forceinline__ double simpson_adaptive(const F& f, double lb, double ub, double ZT, double simps_err,
                                                   int max_depth)
{
    struct Frame {
        double lb, ub, S, fb, fe, fm, left;
        int right_pending;
    };
    Frame st[kMaxDepth];
    int sp = 0;

    const double h0 = ub - lb;
    double fb = 0, fe = 0, fm = 0, fd = 0, S = 0;
    double err = simps_err;
    // The integrand is evaluated at ONE place in the code (it is large, and when it is itself an adaptive integral it
    // contains another copy of this walk): stage 0-2 are the three nodes of the whole interval, 3 and 4 the two new
    // midpoints of the current one.
    int stage = 0;
    for (;;) {
        const double c = (ub + lb) / 2.;
        const double at = (stage == 0) ? lb : (stage == 1) ? ub : (stage == 2) ? c : (stage == 3) ? (lb + c) / 2. : (c + ub) / 2.;
        const double val = f(at) / ZT;
        if (stage < 4) {
            if (stage == 0) fb = val;
            else if (stage == 1) fe = val;
            else if (stage == 2) {
                fm = val;
                S = (h0 / 6) * (fb + 4 * fm + fe);
            } else fd = val;
            stage++;
            continue;
        }
        stage = 3;
        const double fee = val;
        const double h = ub - lb;
        const double Sl = (h / 12) * (fb + 4 * fd + fm);
        const double Sr = (h / 12) * (fm + 4 * fee + fe);
        const double S2 = Sl + Sr;
        if ((max_depth - sp) <= 0 || fabs(S2 - S) <= 15 * err) {
            double ret = S2 + (S2 - S) / 15;
            for (;;) {
                if (sp == 0) return ret;
                Frame& top = st[sp - 1];
                if (top.right_pending) {
                    top.right_pending = 0;
                    top.left = ret;
                    lb = top.lb; ub = top.ub; S = top.S; fb = top.fb; fe = top.fe; fm = top.fm;
                    err = ldexp(simps_err, -sp);
                    break;
                }
                ret = top.left + ret;
                sp--;
            }
        } else {
            Frame& nf = st[sp];
            nf.lb = c; nf.ub = ub; nf.S = Sr; nf.fb = fm; nf.fe = fe; nf.fm = fee; nf.right_pending = 1;
            sp++;
            ub = c; S = Sl; fe = fm; fm = fd;  // descend left: (lb, c), f_beg stays
            err = err / 2;
        }
    }
}

// Inner z-integral used as the integrand of the t-integral (2-D cases).
struct InnerFixed {
    Trial q; double lb_z, ub_z; int n_sz;
    __device__ __forceinline__ double operator()(double t_node) const
    {
        Axis ax{q, t_node, true};
        return simpson_fixed(ax, lb_z, ub_z, n_sz);
    }
};
struct InnerAdaptive {
    Trial q; double lb_z, ub_z, simps_err; int depth_sz;
    __device__ __forceinline__ double operator()(double t_node) const
    {
        Axis ax{q, t_node, true};
        return simpson_adaptive(ax, lb_z, ub_z, ub_z - lb_z, simps_err, depth_sz);
    }
};

// pdf.pxi:104-146
__device__ __forceinline__ double full_pdf(double x, double v, double sv, double a, double z, double sz, double t,
                                           double st, const Opts& o)
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

    const Trial q{x, v, sv, a, o.err};
    if (sz == 0 && st == 0) return pdf_sv(x - t, v, sv, a, z, o.err);
    const double lb_t = t - st / 2., ub_t = t + st / 2.;
    const double lb_z = z - sz / 2., ub_z = z + sz / 2.;
    if (sz == 0 || st == 0) {
        // one axis.  sz == 0: adaptiveSimpsons_1D picks the axis by (ub_t - lb_t) == 0 (integrate.pxi:120), simpson_1D
        // by n_st == 0 (integrate.pxi:17-28); both then integrate "over z" on the empty interval [z, z].
        const bool adapt = o.use_adaptive > 0 || (st == 0 && o.use_adaptive);
        bool over_z;
        double lo, hi;
        int n;
        if (st == 0) {
            over_z = true; lo = lb_z; hi = ub_z; n = o.n_sz;
        } else {
            over_z = adapt ? ((ub_t - lb_t) == 0) : (o.n_st == 0);
            lo = over_z ? z : lb_t; hi = over_z ? z : ub_t; n = o.n_st;
        }
        const Axis ax{q, over_z ? t : z, over_z};
        if (adapt) return simpson_adaptive(ax, lo, hi, (st == 0) ? ub_z - lb_z : (over_z ? 0.0 : ub_t - lb_t), o.simps_err, n);
        return simpson_fixed(ax, lo, hi, n > 0 ? n : 0);
    }
    if (o.use_adaptive) {
        InnerAdaptive in{q, lb_z, ub_z, o.simps_err, o.n_sz};
        return simpson_adaptive(in, lb_t, ub_t, ub_t - lb_t, o.simps_err, o.n_st);
    }
    InnerFixed in{q, lb_z, ub_z, o.n_sz};
    // simpson_2D (integrate.pxi:47-70): same 1-4-2-1 walk over t, each node an inner z-integral
    return simpson_fixed(in, lb_t, ub_t, o.n_st);
}

// pdf.pxi:67-72
__device__ __forceinline__ double prob_ub(double v, double a, double z)
{
    if (v == 0) return z;
    return (exp(-2 * a * z * v) - 1) / (exp(-2 * a * v) - 1);
}

}  // namespace faithful
}  // namespace wfpt
