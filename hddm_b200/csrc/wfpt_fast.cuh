// Restructured ("fast") device evaluation of the WFPT density: same mathematics and the same series term count K
// as the reference, far fewer FP64 instructions.  What changes relative to the operation-for-operation variant:
//
//  * K and the small/large-time choice come from a table lookup on the bit pattern of tt (ktable.h) instead of
//    2 log + 6 sqrt + 2 div per node.  The table is built from the reference's own expressions, so K is the
//    reference's K.
//  * Everything that depends only on the parameter set is computed once per CTA and staged in shared memory: the
//    quadrature nodes z_j (for both response signs) and t_i with the reference's roundings, the drift-variability
//    exponent coefficients A_j = (a z_j sv)^2 - 2 a v z_j, and the large-time sines k*sin(k*pi*z_j).
//  * Everything that depends on (trial, t-node) only is shared by the z-nodes of that t-node: tt, 1/tt, K,
//    q^(k^2) = exp(-k^2 pi^2 tt/2) by repeated multiplication, 1/(2(sv^2 x+1)), the 1/sqrt prefactors.
//  * exp(log(p) + E) is evaluated as p*exp(E) with E folded into the small-time exponents, so a small-time node costs
//    three exp() (k = -1, 0, 1; further terms by the Gaussian ratio recurrence) and a large-time node one.
//  * pdf_sv's NaN for a negative truncated series (log of a negative p, /root/reference/src/pdf.pxi:102) is kept.
//
// Differences from the reference's result are rounding-level (~1e-15 relative per density); tests hold the summed
// log-likelihood to 1e-10 relative.
#pragma once
#include "common.cuh"
#include "ktable.h"
#include "wfpt_faithful.cuh"  // generic fallback for what the table-driven paths do not cover

namespace wfpt {
namespace fast {

constexpr int kMaxNodes = 33;               // nodes per axis held in shared memory (n_st, n_sz <= 32)
constexpr double kInvSqrt2Pi = 0.3989422804014327;
constexpr double kHalfPiSq = 4.934802200544679;  // pi^2 / 2

enum Mode : int {
    kModePoint = 0,      // sz == 0 and st == 0: one node per trial
    kModeFixed = 1,      // fixed-grid Simpson over t and/or z, nodes from shared memory
    kModeAdaptive = 2,   // adaptive Simpson: first level from shared memory, deeper levels by the generic walker
    kModeGeneric = 3,    // anything else: operation-for-operation evaluator
    kModeZero = 4        // parameter set outside the support: density 0 for every trial
};

// Shared-memory tables of one parameter set (one CTA).
struct Tables {
    double zw[2][kMaxNodes];                 // z_j for sign 0 (lower-boundary RT, v, z) and sign 1 (upper: -v, 1-z)
    double zA[2][kMaxNodes];                 // (a z_j sv)^2 - 2 a v' z_j
    double zS[2][kMaxNodes][kMaxKLarge];     // k * sin(k pi z_j), k = 1..kl_max
    double tn[kMaxNodes];                    // t_i
    double wz[2], hz[2];                     // z interval width / step per response sign, as the reference rounds them
    double wt, ht;                           // same along t
    Params prm;                              // the raw parameter set and options (generic fallback only)
    Opts o;
};

// Per-thread (register) view of the parameter set; everything bulky lives in Tables.
struct Set {
    int mode;
    int nz, nt;          // intervals along z / t used by the table paths (0: axis not integrated)
    int sv_nonzero;
    int depth_z, depth_t;  // adaptive recursion budgets (n_sz, n_st)
    double simps_err;
    double t;
    double tlo;          // t - st/2 (validity bound on |x|)
    double v2, sv2, a2, inv_a2;
    const Tables* tab;
};

// Out-of-line call into the operation-for-operation evaluator, so that its explicit Simpson stacks and register
// appetite stay out of the table-driven paths.
static __device__ __noinline__ double generic_full_pdf(double x, const Params& prm, const Opts& o)
{
    return faithful::full_pdf(x, prm.v, prm.sv, prm.a, prm.z, prm.sz, prm.t, prm.st, o);
}

__device__ __forceinline__ Tables& tables()
{
    __shared__ Tables t;
    return t;
}

// (small, K) for this tt: linear scan over the (few) thresholds; the table sits in the constant bank.
__device__ __forceinline__ int klookup(const KTable& kt, double tt)
{
    const long long b = __double_as_longlong(tt);
    int d = kt.desc0;
#pragma unroll 4
    for (int i = 0; i < kt.n; i++) d = (b >= kt.thr[i]) ? kt.desc[i] : d;
    return d;
}

// Per-(trial, t-node) quantities shared by all z-nodes.
struct TNode {
    int alive;       // x' > 0
    int small;
    int K;
    double E0;       // -v^2 x' * c      (E_j = A_j * c + E0)
    double c;        // 1 / (2 (sv^2 x' + 1))
    double nh;       // -1 / (2 tt)      (small-time)
    double pref;     // everything multiplying the series sum
    double q1, q4, q9, q16; // q^(k^2), k = 1..4 (large-time); higher powers are rebuilt on demand
};

__device__ __forceinline__ void make_tnode(TNode& tn, double xp, const Set& s, const KTable& kt)
{
    tn.alive = xp > 0;
    if (!tn.alive) return;
    const double tt = xp / s.a2;  // the reference's tt, bit for bit: K is decided on it
    const int d = klookup(kt, tt);
    tn.small = d >> 8;
    tn.K = d & 0xff;
    const double den = fma(s.sv2, xp, 1.0);
    tn.c = 0.5 / den;
    tn.E0 = -s.v2 * xp * tn.c;
    const double rs = sqrt(2.0 * tn.c);  // 1/sqrt(sv^2 x' + 1)
    if (tn.small) {
        const double inv_tt = 1.0 / tt;
        tn.nh = -0.5 * inv_tt;
        tn.pref = inv_tt * sqrt(inv_tt) * (kInvSqrt2Pi * s.inv_a2) * rs;
    } else {
        const double q = exp(-kHalfPiSq * tt);
        const double q2 = q * q, q4 = q2 * q2, q8 = q4 * q4;
        tn.q1 = q;
        tn.q4 = q4;
        tn.q9 = q8 * q;
        tn.q16 = q8 * q8;
        tn.pref = kPi * s.inv_a2 * rs;
    }
}

// sum_k (w+2k) exp(-(w+2k)^2/(2tt) + E), k = -floor((K-1)/2) .. ceil((K-1)/2)   (pdf.pxi:51-56 with E folded in)
__device__ __forceinline__ double series_small(double w, int K, double nh, double E)
{
    const double u0 = w, u1 = w + 2.0, um = w - 2.0;
    const double e0 = exp(fma(u0 * u0, nh, E));
    const double e1 = exp(fma(u1 * u1, nh, E));
    if (K <= 2) {
        // K == 2: k = 0, 1;  K == 1: k = 0;  K == 0: no term (cannot occur: ks >= 2)
        return K == 2 ? fma(u1, e1, u0 * e0) : (K == 1 ? u0 * e0 : 0.0);
    }
    const double em = exp(fma(um * um, nh, E));
    double sum = fma(u1, e1, fma(u0, e0, um * em));
    if (K > 3 && e0 > 0.0) {
        // e_{k+1}/e_k = rho * H^k and e_{-(k+1)}/e_{-k} = sig * H^k with H = rho * sig = exp(-4/tt)
        const double inv_e0 = 1.0 / e0;
        const double rho = e1 * inv_e0, sig = em * inv_e0, H = rho * sig;
        const int upper = K >> 1;        // ceil((K-1)/2)
        const int lower = (K - 1) >> 1;  // floor((K-1)/2)
        double ek = e1, r = rho;
        for (int k = 2; k <= upper; k++) {
            r *= H;
            ek *= r;
            sum = fma(w + 2.0 * k, ek, sum);
        }
        ek = em;
        r = sig;
        for (int k = 2; k <= lower; k++) {
            r *= H;
            ek *= r;
            sum = fma(w - 2.0 * k, ek, sum);
        }
    }
    return sum;
}

// One quadrature node: pdf_sv(x', v', sv, a, z_j, err) from the staged tables.
__device__ __forceinline__ double node_table(const TNode& tn, const Set& s, int sign, int j)
{
    if (!tn.alive) return 0.0;
    const Tables& T = *s.tab;
    const double E = fma(T.zA[sign][j], tn.c, tn.E0);
    double series;
    if (tn.small) {
        series = series_small(T.zw[sign][j], tn.K, tn.nh, E);
    } else {
        const double* S = T.zS[sign][j];
        double acc = tn.q1 * S[0];
        if (tn.K > 1) acc = fma(tn.q4, S[1], acc);
        if (tn.K > 2) acc = fma(tn.q9, S[2], acc);
        if (tn.K > 3) acc = fma(tn.q16, S[3], acc);
        if (tn.K > 4) {
            const double q2 = tn.q1 * tn.q1;
            double odd = tn.q4 * tn.q4 * tn.q1;  // q^9 = q^(25-16)
            double qk = tn.q16;
            for (int k = 4; k < tn.K; k++) {      // term k+1: q^((k+1)^2) = q^(k^2) * q^(2k+1)
                qk *= odd;
                odd *= q2;
                acc = fma(qk, S[k], acc);
            }
        }
        if (tn.K < 1) acc = 0.0;
        series = acc * exp(E);
    }
    if (s.sv_nonzero && series < 0.0) return CUDART_NAN;  // log(p) of a negative truncated series (pdf.pxi:102)
    return series * tn.pref;
}

// integrate.pxi:30-45 accumulation order, nodes addressed by index.
template <class F>
__device__ __forceinline__ double simpson_idx(const F& f, int n, double width, double h)
{
    double S = f(0);
    double y = 0.0;
    for (int i = 1; i <= n; i++) {
        y = f(i);
        S += (i & 1) ? (4 * y) : (2 * y);
    }
    S = S - y;
    S = S / width;
    return (h * S / 3);
}

// First level of integrate.pxi:72-141 on five equally spaced nodes f0..f4 (already divided by ZT):
// returns true and the value when the reference would stop here.
__device__ __forceinline__ bool adaptive_level1(double f0, double f1, double f2, double f3, double f4, double h,
                                                double simps_err, int max_depth, double* out)
{
    const double S = (h / 6) * (f0 + 4 * f2 + f4);
    const double Sl = (h / 12) * (f0 + 4 * f1 + f2);
    const double Sr = (h / 12) * (f2 + 4 * f3 + f4);
    const double S2 = Sl + Sr;
    *out = S2 + (S2 - S) / 15;
    return (max_depth <= 0) || (fabs(S2 - S) <= 15 * simps_err);
}

// CTA-collective: classify the parameter set and fill the shared tables.  Every thread of the CTA must call it.
__device__ __forceinline__ void prepare(Set& s, const Params& prm, const Opts& o, const KTable& kt)
{
    Tables& T = tables();
    s.tab = &T;
    const double z = prm.z, a = prm.a, t = prm.t, sv = prm.sv;
    double sz = prm.sz, st = prm.st;
    s.t = t;
    s.tlo = t - st / 2.;
    s.v2 = prm.v * prm.v;
    s.sv2 = sv * sv;
    s.a2 = a * a;
    s.inv_a2 = 1.0 / s.a2;
    s.sv_nonzero = (sv != 0);
    s.simps_err = o.simps_err;
    s.depth_z = o.n_sz;
    s.depth_t = o.n_st;
    s.nz = s.nt = 0;
    // pdf.pxi:111-113, the trial-independent part
    const bool bad = (z < 0) || (z > 1) || (a < 0) || (t < 0) || (st < 0) || (sv < 0) || (sz < 0) || (sz > 1) ||
                     (z + sz / 2. > 1) || (z - sz / 2. < 0) || (t - st / 2. < 0);
    if (st < 1e-3) st = 0;
    if (sz < 1e-3) sz = 0;
    int mode;
    if (bad) mode = kModeZero;
    else if (kt.n < 0 || kt.kl_max > kMaxKLarge) mode = kModeGeneric;
    else if (sz == 0 && st == 0) mode = kModePoint;
    else if (o.use_adaptive) mode = kModeAdaptive;
    else {
        const int nz = (sz == 0) ? 0 : o.n_sz, nt = (st == 0) ? 0 : o.n_st;
        // n == 0 on an integrated axis is the reference's division-by-zero corner (Appendix A.12): generic path
        mode = ((sz != 0 && nz == 0) || (st != 0 && nt == 0) || nz >= kMaxNodes || nt >= kMaxNodes) ? kModeGeneric : kModeFixed;
    }
    s.mode = mode;
    if (threadIdx.x == 0) {
        T.prm = prm;
        T.o = o;
    }
    if (mode == kModeZero || mode == kModeGeneric) {
        __syncthreads();
        return;
    }

    // intervals / nodes per axis
    if (mode == kModeFixed) {
        s.nz = (sz == 0) ? 0 : o.n_sz;
        s.nt = (st == 0) ? 0 : o.n_st;
    } else if (mode == kModeAdaptive) {
        s.nz = (sz == 0) ? 0 : 4;
        s.nt = (st == 0) ? 0 : 4;
    }
    const int nzn = s.nz + 1, ntn = s.nt + 1;
    const int nz_div = s.nz, nt_div = s.nt;

    // t axis (sign independent).  Every operation rounded separately, as integrate.pxi:35-36,61 / :87-96 do.
    {
        const double lb = __dsub_rn(t, __ddiv_rn(st, 2.)), ub = __dadd_rn(t, __ddiv_rn(st, 2.));
        const double wt = __dsub_rn(ub, lb);
        const double ht = (mode == kModeFixed && nt_div > 0) ? __ddiv_rn(wt, (double)nt_div) : 0.0;
        if (threadIdx.x == 0) {
            T.wt = wt;
            T.ht = ht;
        }
        for (int i = threadIdx.x; i < ntn; i += blockDim.x) {
            double ti;
            if (ntn == 1) ti = t;
            else if (mode == kModeFixed) ti = __dadd_rn(lb, __dmul_rn(ht, (double)i));
            else {
                const double c = __ddiv_rn(__dadd_rn(ub, lb), 2.);
                ti = (i == 0) ? lb : (i == 4) ? ub : (i == 2) ? c
                     : (i == 1) ? __ddiv_rn(__dadd_rn(lb, c), 2.) : __ddiv_rn(__dadd_rn(c, ub), 2.);
            }
            T.tn[i] = ti;
        }
    }
    // z axis, both response signs
    auto znode = [&](int sign, int j, double* width) {
        const double zz = sign ? __dsub_rn(1., z) : z;   // pdf.pxi:116-118
        const double lb = __dsub_rn(zz, __ddiv_rn(sz, 2.)), ub = __dadd_rn(zz, __ddiv_rn(sz, 2.));
        *width = __dsub_rn(ub, lb);
        if (nzn == 1) return zz;
        if (mode == kModeFixed) {
            const double hz = __ddiv_rn(*width, (double)nz_div);
            return __dadd_rn(lb, __dmul_rn(hz, (double)j));
        }
        const double c = __ddiv_rn(__dadd_rn(ub, lb), 2.);
        return (j == 0) ? lb : (j == 4) ? ub : (j == 2) ? c
               : (j == 1) ? __ddiv_rn(__dadd_rn(lb, c), 2.) : __ddiv_rn(__dadd_rn(c, ub), 2.);
    };
    if (threadIdx.x < 2) {
        const int sign = threadIdx.x;
        double w;
        (void)znode(sign, 0, &w);
        T.wz[sign] = w;
        T.hz[sign] = (mode == kModeFixed && nz_div > 0) ? __ddiv_rn(w, (double)nz_div) : 0.0;
    }
    for (int idx = threadIdx.x; idx < 2 * nzn; idx += blockDim.x) {
        const int sign = idx / nzn, j = idx - sign * nzn;
        double w;
        const double zj = znode(sign, j, &w);
        const double vv = sign ? -prm.v : prm.v;
        T.zw[sign][j] = zj;
        const double azsv = a * zj * sv;
        T.zA[sign][j] = azsv * azsv - 2 * a * vv * zj;
    }
    // k sin(k pi z_j): one (sign, node, k) entry per thread
    for (int idx = threadIdx.x; idx < 2 * nzn * kMaxKLarge; idx += blockDim.x) {
        const int k = idx % kMaxKLarge, r = idx / kMaxKLarge;
        const int sign = r / nzn, j = r - sign * nzn;
        double w;
        const double zj = znode(sign, j, &w);
        T.zS[sign][j][k] = (k < kt.kl_max) ? (double)(k + 1) * sinpi((double)(k + 1) * zj) : 0.0;
    }
    __syncthreads();
}

// Inner z-integral at one t-node: a single node, a fixed grid, or the first adaptive level (ok &= converged).
__device__ __forceinline__ double inner_z(const TNode& tn, const Set& s, int sign, bool* ok)
{
    if (s.nz == 0) return node_table(tn, s, sign, 0);
    const Tables& T = *s.tab;
    const double wz = T.wz[sign];
    if (s.mode == kModeFixed) {
        return simpson_idx([&](int j) { return node_table(tn, s, sign, j); }, s.nz, wz, T.hz[sign]);
    }
    const double f0 = node_table(tn, s, sign, 0) / wz;
    const double f1 = node_table(tn, s, sign, 1) / wz;
    const double f2 = node_table(tn, s, sign, 2) / wz;
    const double f3 = node_table(tn, s, sign, 3) / wz;
    const double f4 = node_table(tn, s, sign, 4) / wz;
    double val;
    *ok &= adaptive_level1(f0, f1, f2, f3, f4, wz, s.simps_err, s.depth_z, &val);
    return val;
}

__device__ __forceinline__ double density(double x, const Set& s, const KTable& kt)
{
    if (s.mode == kModeZero) return 0.0;
    const Tables& T = *s.tab;
    if (s.mode == kModeGeneric) return generic_full_pdf(x, T.prm, T.o);
    const double ax = fabs(x);
    if ((ax - s.tlo) < 0) return 0.0;  // pdf.pxi:112, the trial-dependent part
    const int sign = x > 0;
    bool ok = true;
    double result;
    if (s.nt == 0) {
        TNode tn;
        make_tnode(tn, ax - s.t, s, kt);
        result = inner_z(tn, s, sign, &ok);
    } else if (s.mode == kModeFixed) {
        result = simpson_idx(
            [&](int i) {
                TNode tn;
                make_tnode(tn, ax - T.tn[i], s, kt);
                return inner_z(tn, s, sign, &ok);
            },
            s.nt, T.wt, T.ht);
    } else {
        // kModeAdaptive: first level of the reference's recursion on the staged quarter points
        const double wt = T.wt;
        double g[5];
#pragma unroll 1
        for (int i = 0; i < 5; i++) {
            TNode tn;
            make_tnode(tn, ax - T.tn[i], s, kt);
            g[i] = inner_z(tn, s, sign, &ok) / wt;
        }
        ok &= adaptive_level1(g[0], g[1], g[2], g[3], g[4], wt, s.simps_err, s.depth_t, &result);
    }
    if (ok) return result;
    // some interval needs refinement: walk the full recursion for this trial
    return generic_full_pdf(x, T.prm, T.o);
}

// Per-trial parameter sets (wiener_like_multi, RL-DDM): nothing can be staged per CTA.  The common regression case
// (sz, st below the 1e-3 threshold) is one node evaluated from registers; everything else takes the generic evaluator.
__device__ __forceinline__ double density_once(double x, const Params& prm, const Opts& o, const KTable& kt)
{
    const double z = prm.z, a = prm.a, t = prm.t, sv = prm.sv, sz = prm.sz, st = prm.st;
    if (!(sz < 1e-3 && st < 1e-3) || kt.n < 0 || kt.kl_max > kMaxKLarge) return generic_full_pdf(x, prm, o);
    if ((z < 0) || (z > 1) || (a < 0) || (t < 0) || (st < 0) || (sv < 0) || (sz < 0) || (sz > 1) ||
        ((fabs(x) - (t - st / 2.)) < 0) || (z + sz / 2. > 1) || (z - sz / 2. < 0) || (t - st / 2. < 0))
        return 0.0;
    const int sign = x > 0;
    const double w = sign ? __dsub_rn(1., z) : z;
    const double vv = sign ? -prm.v : prm.v;
    const double xp = fabs(x) - t;
    if (!(xp > 0)) return 0.0;
    const double a2 = a * a, inv_a2 = 1.0 / a2;
    const double tt = xp / a2;
    const int d = klookup(kt, tt);
    const int K = d & 0xff;
    const double c = 0.5 / fma(sv * sv, xp, 1.0);
    const double azsv = a * w * sv;
    const double E = (azsv * azsv - 2 * a * vv * w - prm.v * prm.v * xp) * c;
    const double rs = sqrt(2.0 * c);
    double series, pref;
    if (d >> 8) {
        const double inv_tt = 1.0 / tt;
        series = series_small(w, K, -0.5 * inv_tt, E);
        pref = inv_tt * sqrt(inv_tt) * (kInvSqrt2Pi * inv_a2) * rs;
    } else {
        const double q = exp(-kHalfPiSq * tt), q2 = q * q;
        double odd = q, qk = q, acc = 0.0;
        for (int k = 1; k <= K; k++) {
            acc = fma(qk * (double)k, sinpi((double)k * w), acc);
            odd *= q2;
            qk *= odd;
        }
        series = acc * exp(E);
        pref = kPi * inv_a2 * rs;
    }
    if (sv != 0 && series < 0.0) return CUDART_NAN;
    return series * pref;
}

}  // namespace fast
}  // namespace wfpt
