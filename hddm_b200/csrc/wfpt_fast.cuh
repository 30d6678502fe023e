// Restructured ("fast") device evaluation of the WFPT density: same mathematics and the same series term count K
// as the reference, far fewer FP64 instructions, and independent instruction chains for the FP64 pipe to overlap.
// What changes relative to the operation-for-operation variant (wfpt_faithful.cuh):
//
//  * K and the small/large-time choice come from a table lookup on the bit pattern of tt (ktable.h) instead of
//    2 log + 6 sqrt + 2 div per node.  The table is built from the reference's own expressions, so K is the
//    reference's K.  Each thread remembers the piece it was in and walks from there (RTs are |rt|-sorted by the
//    batching layer, so the walk is usually zero steps).
//  * Everything that depends only on the parameter set is computed once per CTA and staged in shared memory: the
//    quadrature nodes z_j (for both response signs) and t_i with the reference's roundings, per node the exponent
//    coefficient A_j = (a z_j sv)^2 - 2 a v z_j, the small-time lattice u = z_j + 2k and -u^2/2, and the large-time
//    sines k*sin(k*pi*z_j).
//  * Everything that depends on (trial, t-node) only is shared by the z-nodes of that t-node: tt (correctly rounded
//    x'/a^2 from a precomputed reciprocal), tt^-1/2 by rsqrt, K, q^(k^2) = exp(-k^2 pi^2 tt/2) by repeated
//    multiplication, (sv^2 x'+1)^-1/2 by rsqrt.  No division is left in the trial loop.
//  * exp(log(p) + E) is evaluated as p*exp(E) with E folded into the small-time exponents; a small-time node costs
//    three or four exp (k = -1..1 or -1..2; further terms by the Gaussian ratio recurrence), a large-time node one.
//  * exp is a branch-free table-driven kernel (2^(j/64) table in shared memory, degree-5 polynomial, < 1 ulp); the
//    exps of all z-nodes of a t-node are issued as one straight-line batch.  Arguments outside the normal range
//    (result subnormal / overflow / NaN) send the batch to the library exp.
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
#ifndef WFPT_MAX_ZNODES
#define WFPT_MAX_ZNODES kMaxNodes
#endif
constexpr int kMaxZNodes = WFPT_MAX_ZNODES;     // z-nodes held per sign (a translation unit that only stages small z grids sets it lower)
static_assert(kMaxZNodes >= 5 && kMaxZNodes <= kMaxNodes, "the first adaptive level stages five z-nodes");
constexpr double kInvSqrt2Pi = 0.3989422804014327;
constexpr double kHalfPiSq = 4.934802200544679;  // pi^2 / 2
constexpr int kQueue = 1024;                // per-CTA parking lot for trials that need the generic evaluator

enum Mode : int {
    kModePoint = 0,      // sz == 0 and st == 0: one node per trial
    kModeFixed = 1,      // fixed-grid Simpson over t and/or z, nodes from shared memory
    kModeAdaptive = 2,   // adaptive Simpson: first level from shared memory, deeper levels by the generic walker
    kModeGeneric = 3,    // anything else: operation-for-operation evaluator
    kModeZero = 4        // parameter set outside the support: density 0 for every trial
};

// Per (response sign, z-node) constants.
struct alignas(16) NodeTab {   // field order and alignment chosen so that the node loop can fetch pairs (16-byte loads)
    double u[4];              // z + 2k,        k = -1, 0, 1, 2
    double A;                 // (a z sv)^2 - 2 a v' z
    double uh1;               // -z^2/2
    double S[kMaxKLarge];     // k sin(k pi z), k = 1..kl_max
};

// Shared-memory tables of one parameter set (one CTA).
struct Tables {
    NodeTab z[2][kMaxZNodes];  // sign 0: lower-boundary RT (v, z); sign 1: upper (-v, 1-z)
    double tn[kMaxNodes];     // t_i
    double wz[2], hz[2];      // z interval width / step per response sign, as the reference rounds them
    double wt, ht;            // same along t
    double cz[2], ct;         // hz/(3 wz) per sign and ht/(3 wt): the fixed-grid Simpson scale as one multiplier
    double inv_wz[2], inv_wt; // reciprocal widths (first adaptive level)
    double zstep[2];          // spacing of the staged z nodes per sign (lattice factor exp(-2 step / tt))
    double zq[2][4];          // per sign: the exponent of node n is quadratic in n; coefficients of tau and c in its
                              // linear (beta: [2] tau, [3] c) and quadratic (gamma: [0] tau, [1] c) terms
    double wts_t[kMaxNodes];  // Simpson weights 1 4 2 ... 4 1 along t (fixed grid)
    double wts_z[kMaxZNodes];  // same along z
    double exptab[64];        // 2^(j/64)
    long long thr[kMaxPieces];  // series-term table (copy of the launch's KTable)
    int desc[kMaxPieces + 1];   // desc[0] = piece below thr[0]
    Params prm;               // raw parameter set and options (generic fallback only)
    Opts o;
    int qcount;               // trials parked for the generic evaluator (adaptive refinement needed)
    int depth_z, depth_t;     // adaptive recursion budgets (n_sz, n_st)
    int npieces;              // thresholds in the series-term table
    double simps_err;
};

// Per-thread (register) view of the parameter set; everything bulky lives in Tables.
struct Set {
    int mode;
    int nz, nt;            // intervals along z / t used by the table paths (0: axis not integrated)
    int sv_nonzero;
    int kl_max;            // largest large-time K of the table
    int kind;              // which specialised trial evaluator applies (see density)
    int piece;             // piece of the last lookup (walk start)
    int piece_desc;        // its descriptor, and the bit patterns bounding it: lo_b <= bits(tt) < hi_b
    long long lo_b, hi_b;
    double tlo;            // t - st/2 (validity bound on |x|)
    double v2, sv2, a2, inv_a2;
    const Tables* tab;
};

// Out-of-line call into the operation-for-operation evaluator, so that its explicit Simpson stacks and register
// appetite stay out of the table-driven paths.
static __device__ __noinline__ double generic_full_pdf(double x, const Params& prm, const Opts& o)
{
    return faithful::full_pdf(x, prm.v, prm.sv, prm.a, prm.z, prm.sz, prm.t, prm.st, o);
}

// 2^(i/64), i = 0 .. 63, correctly rounded (mpmath, 200 bits): the per-CTA exp table is copied from here -- calling exp2()
// for it in every CTA's prologue was 2.6 % of the headline kernel's time.
__device__ const double kExp2_64[64] = {
    0x1.0000000000000p+0, 0x1.02c9a3e778061p+0, 0x1.059b0d3158574p+0, 0x1.0874518759bc8p+0,
    0x1.0b5586cf9890fp+0, 0x1.0e3ec32d3d1a2p+0, 0x1.11301d0125b51p+0, 0x1.1429aaea92de0p+0,
    0x1.172b83c7d517bp+0, 0x1.1a35beb6fcb75p+0, 0x1.1d4873168b9aap+0, 0x1.2063b88628cd6p+0,
    0x1.2387a6e756238p+0, 0x1.26b4565e27cddp+0, 0x1.29e9df51fdee1p+0, 0x1.2d285a6e4030bp+0,
    0x1.306fe0a31b715p+0, 0x1.33c08b26416ffp+0, 0x1.371a7373aa9cbp+0, 0x1.3a7db34e59ff7p+0,
    0x1.3dea64c123422p+0, 0x1.4160a21f72e2ap+0, 0x1.44e086061892dp+0, 0x1.486a2b5c13cd0p+0,
    0x1.4bfdad5362a27p+0, 0x1.4f9b2769d2ca7p+0, 0x1.5342b569d4f82p+0, 0x1.56f4736b527dap+0,
    0x1.5ab07dd485429p+0, 0x1.5e76f15ad2148p+0, 0x1.6247eb03a5585p+0, 0x1.6623882552225p+0,
    0x1.6a09e667f3bcdp+0, 0x1.6dfb23c651a2fp+0, 0x1.71f75e8ec5f74p+0, 0x1.75feb564267c9p+0,
    0x1.7a11473eb0187p+0, 0x1.7e2f336cf4e62p+0, 0x1.82589994cce13p+0, 0x1.868d99b4492edp+0,
    0x1.8ace5422aa0dbp+0, 0x1.8f1ae99157736p+0, 0x1.93737b0cdc5e5p+0, 0x1.97d829fde4e50p+0,
    0x1.9c49182a3f090p+0, 0x1.a0c667b5de565p+0, 0x1.a5503b23e255dp+0, 0x1.a9e6b5579fdbfp+0,
    0x1.ae89f995ad3adp+0, 0x1.b33a2b84f15fbp+0, 0x1.b7f76f2fb5e47p+0, 0x1.bcc1e904bc1d2p+0,
    0x1.c199bdd85529cp+0, 0x1.c67f12e57d14bp+0, 0x1.cb720dcef9069p+0, 0x1.d072d4a07897cp+0,
    0x1.d5818dcfba487p+0, 0x1.da9e603db3285p+0, 0x1.dfc97337b9b5fp+0, 0x1.e502ee78b3ff6p+0,
    0x1.ea4afa2a490dap+0, 0x1.efa1bee615a27p+0, 0x1.f50765b6e4540p+0, 0x1.fa7c1819e90d8p+0
};

__device__ __forceinline__ Tables& tables()
{
    __shared__ Tables t;
    return t;
}

// Per-CTA parking lot for trials that wait for another evaluator (count in Tables::qcount).  Apart from Tables so that only
// kernels that park anything carry its 8 KB: the fixed-grid lean kernel does not, and runs with that much more L1.
__device__ __forceinline__ double* parking_lot()
{
    __shared__ double q[kQueue];
    return q;
}

// ---- exp ------------------------------------------------------------------------------------------------------------------
// exp(x) = 2^e * 2^(j/64) * exp(r),  k = rint(64 x / ln 2) = 64 e + j,  r = x - k ln2/64, |r| <= 0.0055.
// Branch-free; valid for |x| < 708 (normal results).  Max error 1.24 ulp (checked against mpmath with exact-FMA emulation).
// The coefficients sit in constant memory: as literals the compiler re-materialises each one with two 32-bit moves
// every time register pressure evicts it (16 % of the trial loop's instructions in the profile of the literal form).
__constant__ double kExpC[10] = {
    92.33248261689366,         // 64 / ln 2
    6755399441055744.0,        // 1.5 * 2^52: adding it leaves rint(x * 64/ln2) in the low word
    -0.01083042469326756,      // -(ln2/64) high part (32 bits: kd * hi is exact)
    -2.9815858269852933e-12,   // -(ln2/64) low part
    1.0 / 120.0, 1.0 / 24.0, 1.0 / 6.0, 0.5, 1.0, 0.0};

__device__ __forceinline__ double exp_core(double x, const double* __restrict__ tab)
{
    const double t = fma(x, kExpC[0], kExpC[1]);
    const int ki = __double2loint(t);
    const double kd = t - kExpC[1];
    double r = fma(kd, kExpC[2], x);
    r = fma(kd, kExpC[3], r);
    // exp(r) - 1 = r + r^2 (1/2 + r/6 + r^2 (1/24 + r/120)), Estrin form: the kernel waits on dependent FP64 results far
    // more than on issue slots, and this chain is two steps shorter than Horner's for the same number of operations
    const double r2 = r * r;
    const double ph = fma(r, kExpC[4], kExpC[5]);
    const double pl = fma(r, kExpC[6], kExpC[7]);
    const double p = fma(r2, fma(ph, r2, pl), r);
    const double T = tab[ki & 63];
    const double res = fma(T, p, T);
    // add e = k >> 6 to the exponent field: one mask and one multiply-add (k - j is a multiple of 64, times 2^14 = e << 20)
    int hi;
    asm("mad.lo.s32 %0, %1, 16384, %2;" : "=r"(hi) : "r"(ki & ~63), "r"(__double2hiint(res)));
    return __hiloint2double(hi, __double2loint(res));
}
// |x| < 708 (and not NaN), tested on the high word so that it costs no FP64-pipe slot
__device__ __forceinline__ bool exp_arg_ok(double x) { return (unsigned)(__double2hiint(x) & 0x7fffffff) < 0x40862000u; }
// exp of a non-positive argument whose true value may underflow: 0 below the kernel's range
__device__ __forceinline__ double exp_core_or_zero(double x, const double* __restrict__ tab)
{
    const double v = exp_core(x, tab);
    return exp_arg_ok(x) ? v : 0.0;
}

// M exps at once, written stage by stage: the M dependent chains (8 FP64 results deep each) then sit next to each other in
// the instruction stream and overlap, instead of running one after the other.  Same arithmetic as exp_core.
template <int M>
__device__ __forceinline__ void exp_core_batch(const double (&x)[M], double (&out)[M], const double* __restrict__ tab)
{
    double t[M], r[M], p[M], T[M];
    int ki[M];
#pragma unroll
    for (int i = 0; i < M; i++) t[i] = fma(x[i], kExpC[0], kExpC[1]);
#pragma unroll
    for (int i = 0; i < M; i++) {
        ki[i] = __double2loint(t[i]);
        t[i] = t[i] - kExpC[1];
    }
#pragma unroll
    for (int i = 0; i < M; i++) r[i] = fma(t[i], kExpC[2], x[i]);
#pragma unroll
    for (int i = 0; i < M; i++) {
        r[i] = fma(t[i], kExpC[3], r[i]);
        T[i] = tab[ki[i] & 63];
    }
#pragma unroll
    for (int i = 0; i < M; i++) {
        const double r2 = r[i] * r[i];
        const double ph = fma(r[i], kExpC[4], kExpC[5]);
        const double pl = fma(r[i], kExpC[6], kExpC[7]);
        p[i] = fma(r2, fma(ph, r2, pl), r[i]);
    }
#pragma unroll
    for (int i = 0; i < M; i++) {
        const double res = fma(T[i], p[i], T[i]);
        int hi;
        asm("mad.lo.s32 %0, %1, 16384, %2;" : "=r"(hi) : "r"(ki[i] & ~63), "r"(__double2hiint(res)));
        out[i] = __hiloint2double(hi, __double2loint(res));
    }
}

// 1/sqrt(x) for normal positive x: hardware seed + two Newton steps (straight-line; <= 2 ulp)
__device__ __forceinline__ double rsqrt_fast(double x)
{
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double hx = 0.5 * x;
    double e = fma(-hx * y, y, 0.5);
    y = fma(y, e, y);
    e = fma(-hx * y, y, 0.5);
    y = fma(y, e, y);
    return y;
}

// ---- series-term lookup -----------------------------------------------------------------------------------------------
// (small << 8) | K for this tt.  The thread keeps the piece of its previous lookup with its bounds in registers: RTs are
// |rt|-sorted, so the next tt nearly always falls into the same piece and the lookup is two 64-bit compares; otherwise
// walk to the right piece through the shared-memory copy of the table.
__device__ __forceinline__ int klookup(Set& s, double tt)
{
    const long long b = __double_as_longlong(tt);
    if (b < s.lo_b || b >= s.hi_b) {
        const Tables& T = *s.tab;
        int p = s.piece;
        const int n = T.npieces;
        while (p < n && b >= T.thr[p]) p++;
        while (p > 0 && b < T.thr[p - 1]) p--;
        s.piece = p;
        s.piece_desc = T.desc[p];
        s.lo_b = (p > 0) ? T.thr[p - 1] : (long long)0x8000000000000000ull;
        s.hi_b = (p < n) ? T.thr[p] : (long long)0x7fffffffffffffffull;
    }
    return s.piece_desc;
}

// Scan form for callers without shared tables (per-trial parameter sets).
__device__ __forceinline__ int klookup_scan(const KTable& kt, double tt)
{
    const long long b = __double_as_longlong(tt);
    int d = kt.desc0;
#pragma unroll 4
    for (int i = 0; i < kt.n; i++) d = (b >= kt.thr[i]) ? kt.desc[i] : d;
    return d;
}

// ---- per-(trial, t-node) quantities shared by all z-nodes ---------------------------------------------------------------
struct TNode {
    int alive;       // x' > 0
    int small;
    int K;
    double E0;       // -v^2 x' * c      (E_j = A_j * c + E0)
    double c;        // 1 / (2 (sv^2 x' + 1))
    double inv_tt;   // small-time
    double pref;     // everything multiplying the series sum
    double q1, q4, q9, q16;  // q^(k^2), k = 1..4 (large-time); higher powers are rebuilt on demand
};

__device__ __forceinline__ void make_tnode(TNode& tn, double xp, Set& s)
{
    tn.alive = xp > 0;
    if (!tn.alive) return;
    const Tables& T = *s.tab;
    // tt = x'/a^2, correctly rounded from the reciprocal (quotient + one residual correction): K is decided on it
    const double q0 = xp * s.inv_a2;
    const double tt = fma(fma(-q0, s.a2, xp), s.inv_a2, q0);
    const int d = klookup(s, tt);
    tn.small = d >> 8;
    tn.K = d & 0xff;
    double rs = 1.0;
    tn.c = 0.5;
    if (s.sv_nonzero) {
        rs = rsqrt_fast(fma(s.sv2, xp, 1.0));  // 1/sqrt(sv^2 x' + 1)
        tn.c = 0.5 * rs * rs;
    }
    tn.E0 = -s.v2 * xp * tn.c;
    if (tn.small) {
        const double r = rsqrt_fast(tt);
        tn.inv_tt = r * r;
        tn.pref = tn.inv_tt * r * (kInvSqrt2Pi * s.inv_a2) * rs;  // tt^-3/2 / sqrt(2 pi) / a^2 / sqrt(sv^2 x'+1)
    } else {
        const double qa = -kHalfPiSq * tt;
        const double q = exp_arg_ok(qa) ? exp_core(qa, T.exptab) : exp(qa);
        const double q2 = q * q, q4 = q2 * q2, q8 = q4 * q4;
        // powers beyond this node's K are zeroed here so that the node sum below needs no per-term test
        tn.q1 = (tn.K > 0) ? q : 0.0;
        tn.q4 = (tn.K > 1) ? q4 : 0.0;
        tn.q9 = (tn.K > 2) ? q8 * q : 0.0;
        tn.q16 = (tn.K > 3) ? q8 * q8 : 0.0;
        tn.pref = kPi * s.inv_a2 * rs;
    }
}

// Small-time series with the library exp and every term the reference sums (any K); used when an exponent leaves
// exp_core's range and for K >= 5.   sum_k (w+2k) exp(-(w+2k)^2/(2tt) + E), k = -floor((K-1)/2) .. ceil((K-1)/2)
static __device__ __noinline__ double series_small_slow(double w, int K, double inv_tt, double E)
{
    const int upper = K >> 1, lower = (K - 1) >> 1;
    double sum = 0.0;
    for (int k = -lower; k <= upper; k++) {
        const double u = w + 2.0 * k;
        sum = fma(u, exp(fma(-0.5 * u * u, inv_tt, E)), sum);
    }
    return sum;
}

// N consecutive, equally spaced z-nodes (j0 .. j0+N-1) of one t-node: pdf_sv(x', v', sv, a, z_j, err) from the staged
// tables.  `neg` collects the sign bits of the series sums (pdf_sv turns a negative truncated series into NaN).
//
// Small-time nodes: with e0_n = exp(-z_n^2/(2tt) + E_n) (the k = 0 term, evaluated directly), the neighbouring terms are
//   e(k=+1)/e0 = R_n = exp(-2(z_n+1)/tt),   e(k=-1)/e0 = Q_n = exp(-2(1-z_n)/tt),   e(k=+2)/e(k=+1) = R_n * H,
// H = R_n Q_n = exp(-4/tt), and along equally spaced nodes R_n = R_0 rho^n, Q_n = Q_last rho^(last-n) with
// rho = exp(-2 step/tt).  All factors are <= 1, so nothing overflows, and the dominant term keeps full accuracy:
// a t-node with N z-nodes costs N + 3 exps whatever K is (K <= 4; larger K takes the term-by-term routine).
template <int N>
__device__ __forceinline__ void nodes(const TNode& tn, const Set& s, int sign, int j0, double* __restrict__ y, int& neg)
{
    if (!tn.alive) {
#pragma unroll
        for (int n = 0; n < N; n++) y[n] = 0.0;
        return;
    }
    const Tables& T = *s.tab;
    const NodeTab* Z = &T.z[sign][j0];
    const double* tab = T.exptab;
    double E[N];
#pragma unroll
    for (int n = 0; n < N; n++) E[n] = fma(Z[n].A, tn.c, tn.E0);

    if (tn.small) {
        const int K = tn.K;
        const double tau = tn.inv_tt, m2t = -2.0 * tau;
        bool ok = (K <= 4);
        double e0[N];
        const double z_first = Z[0].u[1], z_last = Z[N - 1].u[1];
        double R, Q, rho;
        if (N == 5) {
            // Five equally spaced nodes (first adaptive level): the exponent of e0_n is quadratic in n,
            // a_n = a_0 + n beta + n^2 gamma, so e0_{n+1} / e0_n = exp(beta + gamma) exp(2 gamma)^n: three exps and a few
            // products instead of five exps.  Six exps in all for the node set (with R, Q, rho).
            const double* zq = T.zq[sign];
            const double gam = fma(zq[1], tn.c, zq[0] * tau), bet = fma(zq[3], tn.c, zq[2] * tau);
            double arg[6], val[6];
            arg[0] = fma(Z[0].uh1, tau, E[0]);
            arg[1] = bet + gam;                                              // a_1 - a_0
            arg[2] = 2.0 * gam;
            arg[3] = fma(m2t, z_first, m2t);                                 // R = exp(-2 (z_0 + 1) / tt)
            arg[4] = fma(-m2t, z_last, m2t);                                 // Q = exp(-2 (1 - z_last) / tt)
            arg[5] = m2t * T.zstep[sign];                                    // rho = exp(-2 step / tt)
            // every e0_n and every ratio between neighbours must be inside the exp kernel's range
            ok &= exp_arg_ok(arg[0]) & exp_arg_ok(fma(4.0, bet, fma(16.0, gam, arg[0])));
            ok &= (fabs(arg[1]) < 700.0) & (fabs(fma(7.0, gam, bet)) < 700.0);
            exp_core_batch<6>(arg, val, tab);
            double ratio = val[1];
            e0[0] = val[0];
#pragma unroll
            for (int n = 1; n < 5; n++) {
                e0[n] = e0[n - 1] * ratio;
                ratio *= val[2];
            }
            R = exp_arg_ok(arg[3]) ? val[3] : 0.0;                           // (true value underflows: 0)
            Q = exp_arg_ok(arg[4]) ? val[4] : 0.0;
            rho = exp_arg_ok(arg[5]) ? val[5] : 0.0;
        } else {
            // the N + 3 exps of the node set as one batch: e0_n, then R, Q and rho
            constexpr int M = N + (N > 1 ? 3 : 2);
            double arg[M], val[M];
#pragma unroll
            for (int n = 0; n < N; n++) {
                arg[n] = fma(Z[n].uh1, tau, E[n]);
                ok &= exp_arg_ok(arg[n]);
            }
            arg[N] = fma(m2t, z_first, m2t);                                 // R = exp(-2 (z_0 + 1) / tt)
            arg[N + 1] = fma(-m2t, z_last, m2t);                             // Q = exp(-2 (1 - z_last) / tt)
            if (N > 1) arg[M - 1] = m2t * T.zstep[sign];                     // rho = exp(-2 step / tt)
            exp_core_batch<M>(arg, val, tab);
#pragma unroll
            for (int n = 0; n < N; n++) e0[n] = val[n];
            R = exp_arg_ok(arg[N]) ? val[N] : 0.0;                           // (true value underflows: 0)
            Q = exp_arg_ok(arg[N + 1]) ? val[N + 1] : 0.0;
            rho = (N > 1) ? (exp_arg_ok(arg[M - 1]) ? val[M - 1] : 0.0) : 1.0;
        }
        if (K < 3) Q = 0.0;                                                  // K == 2 sums k = 0, 1 only
        double Rn[N], Qn[N];
        Rn[0] = R;
        Qn[N - 1] = Q;
#pragma unroll
        for (int n = 1; n < N; n++) {
            Rn[n] = Rn[n - 1] * rho;
            Qn[N - 1 - n] = Qn[N - n] * rho;
        }
        const double H = Rn[0] * Qn[0];  // exp(-4/tt); only the k = 2 term (K >= 4) needs it
#pragma unroll
        for (int n = 0; n < N; n++) {
            double inner = fma(Z[n].u[2], Rn[n], Z[n].u[1]);
            inner = fma(Z[n].u[0], Qn[n], inner);
            if (K >= 4) inner = fma(Z[n].u[3], Rn[n] * Rn[n] * H, inner);
            y[n] = fma(e0[n], inner, 0.0);  // + 0.0: an underflowed e0 times a negative sum must give +0, not -0
        }
        if (!ok) {
#pragma unroll  // (a rolled loop would index y[] dynamically and push the whole array into local memory)
            for (int n = 0; n < N; n++) y[n] = series_small_slow(Z[n].u[1], K, tn.inv_tt, E[n]);
        }
    } else {
        bool ok = true;
        double ex[N];
        if (N == 5) {
            // as above for exp(E_n): E_n = E_0 + n beta + n^2 gamma with the c-parts of beta and gamma only
            const double* zq = T.zq[sign];
            const double gam = zq[1] * tn.c, bet = zq[3] * tn.c;
            double arg[3] = {E[0], bet + gam, 2.0 * gam}, val[3];
            ok &= exp_arg_ok(E[0]) & exp_arg_ok(E[4]);
            ok &= (fabs(arg[1]) < 700.0) & (fabs(fma(7.0, gam, bet)) < 700.0);
            exp_core_batch<3>(arg, val, tab);
            double ratio = val[1];
            ex[0] = val[0];
#pragma unroll
            for (int n = 1; n < 5; n++) {
                ex[n] = ex[n - 1] * ratio;
                ratio *= val[2];
            }
        } else {
#pragma unroll
            for (int n = 0; n < N; n++) ok &= exp_arg_ok(E[n]);
            exp_core_batch<N>(E, ex, tab);
        }
        if (!ok) {
#pragma unroll
            for (int n = 0; n < N; n++) ex[n] = exp(E[n]);
        }
#pragma unroll
        for (int n = 0; n < N; n++) {
            const double* S = Z[n].S;
            double acc = fma(tn.q16, S[3], fma(tn.q9, S[2], fma(tn.q4, S[1], tn.q1 * S[0])));
            if (s.kl_max > 4 && tn.K > 4) {           // first test is uniform over the launch
                const double q2 = tn.q1 * tn.q1;
                double odd = tn.q4 * tn.q4 * tn.q1;  // q^9 = q^(25-16)
                double qk = tn.q16;
                for (int k = 4; k < tn.K; k++) {      // term k+1: q^((k+1)^2) = q^(k^2) * q^(2k+1)
                    qk *= odd;
                    odd *= q2;
                    acc = fma(qk, S[k], acc);
                }
            }
            y[n] = fma(acc, ex[n], 0.0);
        }
    }
#pragma unroll
    for (int n = 0; n < N; n++) {
        y[n] *= tn.pref;                    // pref > 0: the sign is the series sum's
        neg |= __double2hiint(y[n]);
    }
}

// First level of integrate.pxi:72-141 on five equally spaced nodes f0..f4 (already divided by ZT):
// returns true and the value when the reference would stop here.
__device__ __forceinline__ bool adaptive_level1(double f0, double f1, double f2, double f3, double f4, double h,
                                                double simps_err, int max_depth, double* out)
{
    const double h12 = h * (1.0 / 12.0);
    const double S = (2.0 * h12) * (f0 + 4 * f2 + f4);
    const double Sl = h12 * (f0 + 4 * f1 + f2);
    const double Sr = h12 * (f2 + 4 * f3 + f4);
    const double S2 = Sl + Sr;
    *out = fma(S2 - S, 1.0 / 15.0, S2);
    return (max_depth <= 0) || (fabs(S2 - S) <= 15 * simps_err);
}

// CTA-collective: classify the parameter set and fill the shared tables.  Every thread of the CTA must call it.
__device__ __forceinline__ void prepare(Set& s, const Params& prm, const Opts& o, const KTable& kt)
{
    Tables& T = tables();
    s.tab = &T;
    const double z = prm.z, a = prm.a, t = prm.t, sv = prm.sv;
    double sz = prm.sz, st = prm.st;
    s.tlo = t - st / 2.;
    s.v2 = prm.v * prm.v;
    s.sv2 = sv * sv;
    s.a2 = a * a;
    s.inv_a2 = 1.0 / s.a2;
    s.sv_nonzero = (sv != 0);
    s.kl_max = kt.kl_max;
    s.kind = 0;
    s.piece = 0;
    s.piece_desc = 0;
    s.lo_b = 1;   // empty interval: the first lookup walks
    s.hi_b = 0;
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
        mode = ((sz != 0 && nz == 0) || (st != 0 && nt == 0) || nz >= kMaxZNodes || nt >= kMaxNodes) ? kModeGeneric : kModeFixed;
    }
    s.mode = mode;
    if (threadIdx.x == 0) {
        T.prm = prm;
        T.o = o;
        T.qcount = 0;
        T.simps_err = o.simps_err;
        T.depth_z = o.n_sz;
        T.depth_t = o.n_st;
        T.npieces = kt.n;
    }
    if (mode == kModeZero || mode == kModeGeneric) {
        __syncthreads();
        return;
    }
    for (int i = threadIdx.x; i < 64; i += blockDim.x) T.exptab[i] = kExp2_64[i];
    for (int i = threadIdx.x; i <= kt.n; i += blockDim.x) {
        T.desc[i] = (i == 0) ? kt.desc0 : kt.desc[i - 1];
        if (i < kt.n) T.thr[i] = kt.thr[i];
    }

    // intervals / nodes per axis
    if (mode == kModeFixed) {
        s.nz = (sz == 0) ? 0 : o.n_sz;
        s.nt = (st == 0) ? 0 : o.n_st;
    } else if (mode == kModeAdaptive) {
        s.nz = (sz == 0) ? 0 : 4;
        s.nt = (st == 0) ? 0 : 4;
    }
    if (mode == kModeFixed) s.kind = (s.nz == 0) ? 0 : (s.nz == 2 ? 1 : 2);
    else if (mode == kModeAdaptive) s.kind = (s.nz > 0) ? (s.nt > 0 ? 3 : 4) : 5;
    else s.kind = 0;
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
            T.ct = (nt_div > 0) ? ht / (3.0 * wt) : 1.0;
            T.inv_wt = 1.0 / wt;
        }
        for (int i = threadIdx.x; i < ntn; i += blockDim.x) {
            // integrate.pxi:38-43: 4 on odd nodes, 2 on even ones, then "S - y" takes one f(b) back: the last node weighs
            // 1 for an even number of intervals and 3 for an odd one (the reference's "n even" asserts are commented out)
            T.wts_t[i] = (i == 0) ? 1.0 : (i == ntn - 1) ? ((i & 1) ? 3.0 : 1.0) : ((i & 1) ? 4.0 : 2.0);
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
        const double hz = (mode == kModeFixed && nz_div > 0) ? __ddiv_rn(w, (double)nz_div) : 0.0;
        T.hz[sign] = hz;
        T.zstep[sign] = (nzn > 1) ? w / (double)(nzn - 1) : 0.0;
        {
            const double h = T.zstep[sign], z0 = znode(sign, 0, &w), asv = a * sv, vv = sign ? -prm.v : prm.v;
            T.zq[sign][0] = -0.5 * h * h;
            T.zq[sign][1] = (asv * h) * (asv * h);
            T.zq[sign][2] = -z0 * h;
            T.zq[sign][3] = 2.0 * asv * asv * z0 * h - 2.0 * a * vv * h;
        }
        T.cz[sign] = hz / (3.0 * w);
        T.inv_wz[sign] = 1.0 / w;
    }
    // one (sign, node, slot) entry per thread: slots 0..3 = lattice k = -1..2, slot 4 = A, slots 5.. = sines
    constexpr int kSlots = 5 + kMaxKLarge;
    for (int j = threadIdx.x; j < nzn; j += blockDim.x)
        T.wts_z[j] = (j == 0) ? 1.0 : (j == nzn - 1) ? ((j & 1) ? 3.0 : 1.0) : ((j & 1) ? 4.0 : 2.0);
    for (int idx = threadIdx.x; idx < 2 * nzn * kSlots; idx += blockDim.x) {
        const int slot = idx % kSlots, r = idx / kSlots;
        const int sign = r / nzn, j = r - sign * nzn;
        double w;
        const double zj = znode(sign, j, &w);
        NodeTab& Z = T.z[sign][j];
        if (slot < 4) {
            const double u = zj + 2.0 * (slot - 1);
            Z.u[slot] = u;
            if (slot == 1) Z.uh1 = -0.5 * u * u;
        } else if (slot == 4) {
            const double vv = sign ? -prm.v : prm.v;
            const double azsv = a * zj * sv;
            Z.A = azsv * azsv - 2 * a * vv * zj;
        } else {
            const int k = slot - 4;  // 1..kMaxKLarge
            Z.S[k - 1] = (k <= kt.kl_max) ? (double)k * sinpi((double)k * zj) : 0.0;
        }
    }
    __syncthreads();
}

// Inner z-integral at one t-node.  NZN: staged z-nodes (1: no z-integration, 3: fixed grid n_sz = 2, 5: first adaptive
// level, 0: fixed grid of any size, one node at a time).  ok &= "the adaptive recursion would stop here".
template <int NZN>
__device__ __forceinline__ double inner_t(const TNode& tn, const Set& s, int sign, bool* ok, int& neg)
{
    const Tables& T = *s.tab;
    if (NZN == 1) {
        double y;
        nodes<1>(tn, s, sign, 0, &y, neg);
        return y;
    }
    if (NZN == 5) {
        double y[5];
        nodes<5>(tn, s, sign, 0, y, neg);
        const double iw = T.inv_wz[sign];
        double val;
        *ok &= adaptive_level1(y[0] * iw, y[1] * iw, y[2] * iw, y[3] * iw, y[4] * iw, T.wz[sign], T.simps_err,
                               T.depth_z, &val);
        return val;
    }
    // integrate.pxi:30-45: S = f(0) + 4 f(odd) + 2 f(even) - f(n), then (h/3) * S / width
    double S;
    if (NZN == 3) {
        double y[3];
        nodes<3>(tn, s, sign, 0, y, neg);
        S = fma(4.0, y[1], y[0]) + y[2];
    } else {
        double y;
        S = 0.0;
#pragma unroll 1
        for (int j = 0; j <= s.nz; j++) {
            nodes<1>(tn, s, sign, j, &y, neg);
            S = fma(T.wts_z[j], y, S);
        }
    }
    return S * T.cz[sign];
}

template <int NZN, bool ADAPT_T>
__device__ __forceinline__ double density_t(double ax, int sign, Set& s, bool* ok, int& neg)
{
    const Tables& T = *s.tab;
    TNode tn;
    if (ADAPT_T) {
        // first level of the reference's recursion over t on the staged quarter points
        double g[5];
#pragma unroll 1
        for (int i = 0; i < 5; i++) {
            make_tnode(tn, ax - T.tn[i], s);
            g[i] = inner_t<NZN>(tn, s, sign, ok, neg) * T.inv_wt;
        }
        double result;
        *ok &= adaptive_level1(g[0], g[1], g[2], g[3], g[4], T.wt, T.simps_err, T.depth_t, &result);
        return result;
    }
    double S = 0.0;
#pragma unroll 1
    for (int i = 0; i <= s.nt; i++) {
        make_tnode(tn, ax - T.tn[i], s);  // tn[0] == t when the t axis is not integrated
        S = fma(T.wts_t[i], inner_t<NZN>(tn, s, sign, ok, neg), S);
    }
    return S * T.ct;  // ct == 1 when the t axis is not integrated
}

// Density of one trial from the staged tables.  *defer is set (and the return value is meaningless) when the
// reference's adaptive recursion would go below the first level for this trial: the caller evaluates those trials
// with density_generic, preferably all together at the end (a CTA-wide pass instead of one diverged lane at a time).
// s.kind (uniform over the CTA) picks the evaluator specialised for this parameter set's integration layout.
template <bool ADAPT>
__device__ __forceinline__ double density(double x, Set& s, bool* defer)
{
    *defer = false;
    if (s.mode == kModeZero) return 0.0;
    // (kModeGeneric never gets here: the kernels send such sets to density_generic, see whole_set_generic)
    const double ax = fabs(x);
    if ((ax - s.tlo) < 0) return 0.0;  // pdf.pxi:112, the trial-dependent part
    const int sign = x > 0;
    bool ok = true;
    int neg = 0;
    double result;
    const int kind = s.kind;
    if (kind == 0) result = density_t<1, false>(ax, sign, s, &ok, neg);            // no z-integration, fixed or no t-grid
    else if (ADAPT) {
        if (kind == 3) result = density_t<5, true>(ax, sign, s, &ok, neg);         // adaptive over z and t
        else if (kind == 4) result = density_t<5, false>(ax, sign, s, &ok, neg);   // adaptive over z only
        else result = density_t<1, true>(ax, sign, s, &ok, neg);                   // adaptive over t only
    } else {
        if (kind == 1) result = density_t<3, false>(ax, sign, s, &ok, neg);        // fixed grid, n_sz = 2 (any n_st)
        else result = density_t<0, false>(ax, sign, s, &ok, neg);                  // fixed grid, other n_sz
    }
    // a negative truncated series anywhere makes pdf_sv's log(p) NaN (pdf.pxi:102); pdf (sv == 0) lets it through
    if (s.sv_nonzero && neg < 0) result = CUDART_NAN;
    *defer = !ok;
    return result;
}

// ---- adaptive recursion below the first level --------------------------------------------------------------------------
// One pdf_sv node at an arbitrary (x', z): same table-driven K, same lattice identities and exp kernel as the staged
// path, with the per-node constants computed on the spot (the refinement nodes are not on the staged grid).
struct AdhocCtx {
    double v, v2, sv2, asv2, a, a2, inv_a2;  // v already flipped for upper-boundary RTs; asv2 = (a sv)^2
    int sv_nonzero;
    const Tables* T;
};

__device__ __forceinline__ double pdf_sv_adhoc(const AdhocCtx& c, double xp, double z)
{
    if (!(xp > 0)) return 0.0;
    const Tables& T = *c.T;
    const double q0 = xp * c.inv_a2;
    const double tt = fma(fma(-q0, c.a2, xp), c.inv_a2, q0);
    const long long b = __double_as_longlong(tt);
    int d = T.desc[0];
    for (int i = 0; i < T.npieces; i++) d = (b >= T.thr[i]) ? T.desc[i + 1] : d;
    const int K = d & 0xff;
    double rs = 1.0, cc = 0.5;
    if (c.sv_nonzero) {
        rs = rsqrt_fast(fma(c.sv2, xp, 1.0));
        cc = 0.5 * rs * rs;
    }
    const double E = (fma(c.asv2 * z, z, -2.0 * c.a * c.v * z) - c.v2 * xp) * cc;
    const double* tab = T.exptab;
    double series, pref;
    if (d >> 8) {
        const double r = rsqrt_fast(tt), tau = r * r, m2t = -2.0 * tau;
        pref = tau * r * (kInvSqrt2Pi * c.inv_a2) * rs;
        const double a0 = fma(-0.5 * z * z, tau, E);
        if (K <= 4 && exp_arg_ok(a0)) {
            const double e0 = exp_core(a0, tab);
            const double R = exp_core_or_zero(fma(m2t, z, m2t), tab);
            const double Q = (K >= 3) ? exp_core_or_zero(fma(-m2t, z, m2t), tab) : 0.0;
            double inner = fma(z + 2.0, R, z);
            inner = fma(z - 2.0, Q, inner);
            if (K >= 4) inner = fma(z + 4.0, R * R * (R * Q), inner);
            series = fma(e0, inner, 0.0);
        } else {
            series = series_small_slow(z, K, tau, E);
        }
    } else {
        const double qa = -kHalfPiSq * tt;
        const double q = exp_arg_ok(qa) ? exp_core(qa, tab) : exp(qa);
        const double q2 = q * q;
        double s1, c1;
        sincospi(z, &s1, &c1);
        // k sin(k pi z) by the Chebyshev recurrence, q^(k^2) by running products
        const double two_c = 2.0 * c1;
        double sk_prev = 0.0, sk = s1, odd = q, qk = q, acc = 0.0;
        for (int k = 1; k <= K; k++) {
            acc = fma(qk * (double)k, sk, acc);
            const double nxt = fma(two_c, sk, -sk_prev);
            sk_prev = sk;
            sk = nxt;
            odd *= q2;
            qk *= odd;
        }
        series = fma(acc, exp_arg_ok(E) ? exp_core(E, tab) : exp(E), 0.0);
        pref = kPi * c.inv_a2 * rs;
    }
    if (c.sv_nonzero && series < 0.0) return CUDART_NAN;  // pdf.pxi:102
    return series * pref;
}

struct AdhocAxis {  // integrand along one axis (cf. faithful::Axis)
    AdhocCtx c;
    double x, fixed;
    bool over_z;
    __device__ __forceinline__ double operator()(double s) const
    {
        return over_z ? pdf_sv_adhoc(c, x - fixed, s) : pdf_sv_adhoc(c, x - s, fixed);
    }
};
// Adaptive Simpson along one axis, node by node (one copy in the kernel: every caller goes through here).
static __device__ __noinline__ double adhoc_axis_integral(const AdhocAxis& ax, double lo, double hi, double width,
                                                          double simps_err, int depth)
{
    return faithful::simpson_adaptive(ax, lo, hi, width, simps_err, depth);
}
struct AdhocInner {  // adaptive z-integral as the integrand of the t-integral (integrate.pxi:143-206)
    AdhocCtx c;
    double x, lb_z, ub_z, simps_err;
    int depth_sz;
    __device__ __forceinline__ double operator()(double t_node) const
    {
        const AdhocAxis ax{c, x, t_node, true};
        return adhoc_axis_integral(ax, lb_z, ub_z, ub_z - lb_z, simps_err, depth_sz);
    }
};

// full_pdf's adaptive branches (pdf.pxi:127-146) with the table-driven node evaluator; the recursion is the
// explicit-stack walk of wfpt_faithful.cuh, so interval splitting and summation order are the reference's.
static __device__ __noinline__ double adaptive_full_pdf(double x, const Tables* Tp)
{
    const Tables& T = *Tp;
    const Params& p = T.prm;
    double v = p.v, z = p.z, sz = p.sz, st = p.st;
    const double t = p.t;
    if ((fabs(x) - (t - st / 2.)) < 0) return 0.0;
    if (x > 0) {
        v = -v;
        z = __dsub_rn(1., z);
    }
    x = fabs(x);
    if (st < 1e-3) st = 0;
    if (sz < 1e-3) sz = 0;
    AdhocCtx c;
    c.v = v; c.v2 = v * v; c.sv2 = p.sv * p.sv; c.a = p.a; c.a2 = p.a * p.a; c.inv_a2 = 1.0 / c.a2;
    c.asv2 = c.a2 * c.sv2; c.sv_nonzero = (p.sv != 0); c.T = Tp;
    const double lb_z = __dsub_rn(z, __ddiv_rn(sz, 2.)), ub_z = __dadd_rn(z, __ddiv_rn(sz, 2.));
    const double lb_t = __dsub_rn(t, __ddiv_rn(st, 2.)), ub_t = __dadd_rn(t, __ddiv_rn(st, 2.));
    if (sz == 0 || st == 0) {   // one axis: along t at fixed z, or along z at fixed t
        const bool over_z = (sz != 0);
        const AdhocAxis ax{c, x, over_z ? t : z, over_z};
        const double lo = over_z ? lb_z : lb_t, hi = over_z ? ub_z : ub_t;
        return adhoc_axis_integral(ax, lo, hi, __dsub_rn(hi, lo), T.o.simps_err, over_z ? T.o.n_sz : T.o.n_st);
    }
    AdhocInner in{c, x, lb_z, ub_z, T.o.simps_err, T.o.n_sz};
    return faithful::simpson_adaptive(in, lb_t, ub_t, __dsub_rn(ub_t, lb_t), T.o.simps_err, T.o.n_st);
}

// full_pdf (pdf.pxi:104-146) for a parameter set of its own per trial (wiener_like_multi, the RL-DDM flavours with sz or st
// above the 1e-3 thresholds): nothing can be staged per CTA, but every node still goes through the table-driven evaluator
// above instead of 2 log + 6 sqrt + 2 div + K library exps.  Tp needs exptab and the series-term table only
// (prepare_series_only).  Branch structure as the reference's, node positions of the fixed grids rounded separately.
template <class F>
__device__ __forceinline__ double adhoc_simpson_fixed(const F& f, double lb, double ub, int n)
{
    const double h = __ddiv_rn(__dsub_rn(ub, lb), (double)n);
    double S = 0, y = 0;
    for (int i = 0; i <= n; i++) {   // integrate.pxi:30-45: 1, 4, 2, ..., then one f(b) taken back
        y = f(i == 0 ? lb : __dadd_rn(lb, __dmul_rn(h, (double)i)));
        S = (i == 0) ? y : S + ((i & 1) ? (4 * y) : (2 * y));
    }
    S = S - y;
    S = S / __dsub_rn(ub, lb);
    return (h * S / 3);
}
struct AdhocInnerFixed {   // fixed-grid z-integral as the integrand of the t-integral (integrate.pxi:47-70)
    AdhocCtx c;
    double x, lb_z, ub_z;
    int n_sz;
    __device__ __forceinline__ double operator()(double t_node) const
    {
        const AdhocAxis ax{c, x, t_node, true};
        return adhoc_simpson_fixed(ax, lb_z, ub_z, n_sz);
    }
};

static __device__ __noinline__ double full_pdf_adhoc(double x, const Params& p, const Opts& o, const Tables* Tp)
{
    double v = p.v, z = p.z, sz = p.sz, st = p.st;
    const double t = p.t, a = p.a, sv = p.sv;
    if ((z < 0) || (z > 1) || (a < 0) || (t < 0) || (st < 0) || (sv < 0) || (sz < 0) || (sz > 1) ||
        ((fabs(x) - (t - st / 2.)) < 0) || (z + sz / 2. > 1) || (z - sz / 2. < 0) || (t - st / 2. < 0))
        return 0.0;
    if (x > 0) {
        v = -v;
        z = __dsub_rn(1., z);
    }
    x = fabs(x);
    if (st < 1e-3) st = 0;
    if (sz < 1e-3) sz = 0;
    AdhocCtx c;
    c.v = v; c.v2 = v * v; c.sv2 = sv * sv; c.a = a; c.a2 = a * a; c.inv_a2 = 1.0 / c.a2;
    c.asv2 = c.a2 * c.sv2; c.sv_nonzero = (sv != 0); c.T = Tp;
    if (sz == 0 && st == 0) return pdf_sv_adhoc(c, x - t, z);
    const double lb_z = __dsub_rn(z, __ddiv_rn(sz, 2.)), ub_z = __dadd_rn(z, __ddiv_rn(sz, 2.));
    const double lb_t = __dsub_rn(t, __ddiv_rn(st, 2.)), ub_t = __dadd_rn(t, __ddiv_rn(st, 2.));
    if (sz == 0 || st == 0) {
        // one axis (the reference's choice of axis and interval, see faithful::full_pdf)
        const bool adapt = o.use_adaptive != 0;
        bool over_z;
        double lo, hi;
        int n;
        if (st == 0) {
            over_z = true; lo = lb_z; hi = ub_z; n = o.n_sz;
        } else {
            over_z = adapt ? (__dsub_rn(ub_t, lb_t) == 0) : (o.n_st == 0);
            lo = over_z ? z : lb_t; hi = over_z ? z : ub_t; n = o.n_st;
        }
        const AdhocAxis ax{c, x, over_z ? t : z, over_z};
        if (adapt)
            return adhoc_axis_integral(ax, lo, hi, (st == 0) ? __dsub_rn(ub_z, lb_z) : (over_z ? 0.0 : __dsub_rn(ub_t, lb_t)),
                                       o.simps_err, n);
        return adhoc_simpson_fixed(ax, lo, hi, n > 0 ? n : 0);
    }
    if (o.use_adaptive) {
        const AdhocInner in{c, x, lb_z, ub_z, o.simps_err, o.n_sz};
        return faithful::simpson_adaptive(in, lb_t, ub_t, __dsub_rn(ub_t, lb_t), o.simps_err, o.n_st);
    }
    const AdhocInnerFixed in{c, x, lb_z, ub_z, o.n_sz};
    return adhoc_simpson_fixed(in, lb_t, ub_t, o.n_st);
}

// CTA-collective: the part of prepare() the per-trial evaluator needs (exp table, series-term table).
__device__ __forceinline__ void prepare_series_only(const KTable& kt)
{
    Tables& T = tables();
    if (kt.n < 0) return;
    for (int i = threadIdx.x; i < 64; i += blockDim.x) T.exptab[i] = kExp2_64[i];
    for (int i = threadIdx.x; i <= kt.n; i += blockDim.x) {
        T.desc[i] = (i == 0) ? kt.desc0 : kt.desc[i - 1];
        if (i < kt.n) T.thr[i] = kt.thr[i];
    }
    if (threadIdx.x == 0) T.npieces = kt.n;
    __syncthreads();
}

// Refinement along t with the staged z-nodes.  Nearly every trial that fails the first-level test does so on the t axis
// (an RT close to t +- st/2, where the integrand switches on), while each inner z-integral still stops at its first
// level.  Such a trial needs the reference's walk over t only, and every new t-node costs what a staged one costs:
// the inner integral at an arbitrary t-node is inner_t on the staged z lattice.  When one inner integral does ask for
// refinement, that integral alone is redone node by node.
//
// The reference's walk over t for a recursion budget of 2 (what HDDM passes, likelihoods.py:105-113), written level by
// level: the five nodes of the first level, the four of the second, and -- for each half that asks for it -- the four
// of the third, after which the budget is spent and the half is accepted as it is.  The 17 candidate nodes are visited
// in one loop with one evaluation site, so the lanes of a warp (one parked trial each) stay in step: they differ only
// in whether they take part in the last eight iterations.  Sums are formed in the reference's order
// (left + right at every level).
template <int NZN>
__device__ __forceinline__ double refine_t_depth2(double ax, int sign, Set& s, const AdhocCtx& c, int* neg)
{
    const Tables& T = *s.tab;
    double t[17], g[17];
    t[0] = T.tn[0]; t[4] = T.tn[1]; t[8] = T.tn[2]; t[12] = T.tn[3]; t[16] = T.tn[4];
#pragma unroll
    for (int j = 2; j < 16; j += 4) t[j] = (t[j - 2] + t[j + 2]) / 2.;
#pragma unroll
    for (int j = 1; j < 16; j += 2) t[j] = (t[j - 1] + t[j + 1]) / 2.;
    double part[4] = {0.0, 0.0, 0.0, 0.0};  // level-2 Simpson pieces per quarter: [0,4] [4,8] [8,12] [12,16]
    double half_ret[2] = {0.0, 0.0};
    bool deeper[2] = {false, false};
#pragma unroll 1
    for (int it = 0; it < 17; it++) {
        // it 0..4 -> j = 4 it;  5..8 -> j = 4 (it - 5) + 2;  9..16 -> j = 2 (it - 9) + 1
        const int j = (it < 5) ? 4 * it : (it < 9) ? 4 * (it - 5) + 2 : 2 * (it - 9) + 1;
        if (it == 5) {
            // first level again: a trial parked for an inner integral alone may pass it with the refined values
            double r1;
            if (adaptive_level1(g[0], g[4], g[8], g[12], g[16], T.wt, T.simps_err, T.depth_t, &r1)) return r1;
        }
        if (it == 9) {
            // second level (integrate.pxi:99-112 with bottom = 1, simps_err / 2) on both halves
            const double H = t[16] - t[0], h = t[8] - t[0], h2 = t[16] - t[8];
            const double S_l = (H / 12) * (g[0] + 4 * g[4] + g[8]);    // Sleft / Sright of the first level
            const double S_r = (H / 12) * (g[8] + 4 * g[12] + g[16]);
            part[0] = (h / 12) * (g[0] + 4 * g[2] + g[4]);
            part[1] = (h / 12) * (g[4] + 4 * g[6] + g[8]);
            part[2] = (h2 / 12) * (g[8] + 4 * g[10] + g[12]);
            part[3] = (h2 / 12) * (g[12] + 4 * g[14] + g[16]);
            const double S2l = part[0] + part[1], S2r = part[2] + part[3];
            const double e2 = 15 * (T.simps_err / 2);
            deeper[0] = !(fabs(S2l - S_l) <= e2);
            deeper[1] = !(fabs(S2r - S_r) <= e2);
            half_ret[0] = S2l + (S2l - S_l) / 15;
            half_ret[1] = S2r + (S2r - S_r) / 15;
            if (!deeper[0] && !deeper[1]) break;
        }
        if (it >= 9 && !deeper[(it - 9) >> 2]) continue;
        TNode tn;
        make_tnode(tn, ax - t[j], s);
        bool inner_ok = true;
        double val = inner_t<NZN>(tn, s, sign, &inner_ok, *neg);
        if (NZN == 5 && !inner_ok) {
            // this node's z-integral wants more than its first level: walk it node by node (integrate.pxi:114-141)
            const AdhocAxis az{c, ax, t[j], true};
            val = adhoc_axis_integral(az, T.z[sign][0].u[1], T.z[sign][4].u[1], T.wz[sign], T.simps_err, T.depth_z);
        }
        g[j] = val * T.inv_wt;
    }
    // third level (bottom = 0: accepted whatever the difference), quarter by quarter
#pragma unroll
    for (int hlf = 0; hlf < 2; hlf++) {
        if (!deeper[hlf]) continue;
        double sum = 0.0;
#pragma unroll
        for (int q = 0; q < 2; q++) {
            const int j0 = 8 * hlf + 4 * q;                          // quarter [t[j0], t[j0 + 4]]
            const double h = t[j0 + 4] - t[j0];
            const double Sl = (h / 12) * (g[j0] + 4 * g[j0 + 1] + g[j0 + 2]);
            const double Sr = (h / 12) * (g[j0 + 2] + 4 * g[j0 + 3] + g[j0 + 4]);
            const double S2 = Sl + Sr;
            const double ret = S2 + (S2 - part[2 * hlf + q]) / 15;
            sum = q ? sum + ret : ret;
        }
        half_ret[hlf] = sum;
    }
    return half_ret[0] + half_ret[1];
}

// Per-thread view of the staged parameter set rebuilt from the shared tables (the slow paths keep their own copy so that
// the caller's stays in registers).
__device__ __forceinline__ void rebuild_set(Set& s, const Tables* Tp, int kind, int kl_max)
{
    const Params& p = Tp->prm;
    s.mode = kModeAdaptive;
    s.nz = (kind == 3) ? 4 : 0;
    s.nt = 4;
    s.sv_nonzero = (p.sv != 0);
    s.kl_max = kl_max;
    s.kind = kind;
    s.piece = 0;
    s.piece_desc = 0;
    s.lo_b = 1;
    s.hi_b = 0;
    s.tlo = Tp->tn[0];
    s.v2 = p.v * p.v;
    s.sv2 = p.sv * p.sv;
    s.a2 = p.a * p.a;
    s.inv_a2 = 1.0 / s.a2;
    s.tab = Tp;
}

// The reference's adaptive Simpson walk for a budget of 2 on values that are already there: g[j] = f(pos[j]) / ZT on the
// 17 dyadic points of the interval (0, 4, .. 16 first level; 2, 6, .. second; odd third).  Reads only the points the
// walk visits and reports them in *used (bit j), so that callers may fill the others speculatively or not at all.
__device__ __forceinline__ double walk_depth2(const double* g, const double* pos, double width, double simps_err,
                                              int budget, unsigned* used)
{
    double result;
    *used = 0x11111u;
    if (adaptive_level1(g[0], g[4], g[8], g[12], g[16], width, simps_err, budget, &result)) return result;
    *used |= 0x4444u;
    const double H = pos[16] - pos[0], h = pos[8] - pos[0], h2 = pos[16] - pos[8];
    const double S_l = (H / 12) * (g[0] + 4 * g[4] + g[8]);    // Sleft / Sright of the first level
    const double S_r = (H / 12) * (g[8] + 4 * g[12] + g[16]);
    const double part0 = (h / 12) * (g[0] + 4 * g[2] + g[4]);
    const double part1 = (h / 12) * (g[4] + 4 * g[6] + g[8]);
    const double part2 = (h2 / 12) * (g[8] + 4 * g[10] + g[12]);
    const double part3 = (h2 / 12) * (g[12] + 4 * g[14] + g[16]);
    const double S2l = part0 + part1, S2r = part2 + part3;
    const double e2 = 15 * (simps_err / 2);
    double ret0 = S2l + (S2l - S_l) / 15, ret1 = S2r + (S2r - S_r) / 15;
    auto quarter = [&](int j0, double parent) {                // budget spent: accepted whatever the difference
        const double hq = pos[j0 + 4] - pos[j0];
        const double Sl = (hq / 12) * (g[j0] + 4 * g[j0 + 1] + g[j0 + 2]);
        const double Sr = (hq / 12) * (g[j0 + 2] + 4 * g[j0 + 3] + g[j0 + 4]);
        const double S2 = Sl + Sr;
        return S2 + (S2 - parent) / 15;
    };
    if (!(fabs(S2l - S_l) <= e2)) {
        *used |= 0xaau;
        const double q0 = quarter(0, part0);
        ret0 = q0 + quarter(4, part1);
    }
    if (!(fabs(S2r - S_r) <= e2)) {
        *used |= 0xaa00u;
        const double q0 = quarter(8, part2);
        ret1 = q0 + quarter(12, part3);
    }
    return ret0 + ret1;
}

__device__ __forceinline__ void dyadic17(double* pos, double p0, double p4, double p8, double p12, double p16)
{
    pos[0] = p0; pos[4] = p4; pos[8] = p8; pos[12] = p12; pos[16] = p16;
#pragma unroll
    for (int j = 2; j < 16; j += 4) pos[j] = (pos[j - 2] + pos[j + 2]) / 2.;
#pragma unroll
    for (int j = 1; j < 16; j += 2) pos[j] = (pos[j - 1] + pos[j + 1]) / 2.;
}

// The same walk, one WARP per trial: lane j < 17 evaluates t-node j (all 17 candidates at once, whether or not the
// walk will ask for them), the decisions are then taken on the shared values.  A thread working through a parked trial
// alone needs 9 to 17 node evaluations one after the other (tens of microseconds); this takes the time of one, which is
// what counts when a CTA has only a handful of parked trials (one sweep over the nodes of a hierarchical model, a single
// wiener_like call).  A t-node whose z-integral wants refinement gets the same treatment along z, the whole warp
// working on that node.  Nodes the walk does not visit contribute neither values nor NaN flags.  Every lane of the
// warp must call it with the same x; the result is valid in every lane.
static __device__ __noinline__ double refine_coop(double x, const Tables* Tp, int kind, int kl_max)
{
    Tables& T = *const_cast<Tables*>(Tp);
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5;
    // scratch of the walk (apart from Tables: only kernels that call this routine carry it)
    __shared__ double coop_g[4][18];  // node values of the trial each warp works on
    __shared__ double coop_z[4][18];  // ... and of the z-integral being refined
    __shared__ int coop_f[4][18];     // bit 0: inner integral wants refinement, bit 1: negative series seen
    double* g = coop_g[w];
    double* gz = coop_z[w];
    int* fl = coop_f[w];
    const double ax = fabs(x);
    const int sign = x > 0;
    Set s;
    rebuild_set(s, Tp, kind, kl_max);
    const Params& p = T.prm;
    AdhocCtx c;
    c.v = sign ? -p.v : p.v; c.v2 = s.v2; c.sv2 = s.sv2; c.a = p.a; c.a2 = s.a2; c.inv_a2 = s.inv_a2;
    c.asv2 = c.a2 * c.sv2; c.sv_nonzero = s.sv_nonzero; c.T = Tp;
    double t[17];
    dyadic17(t, T.tn[0], T.tn[1], T.tn[2], T.tn[3], T.tn[4]);
    double tj = t[0];
#pragma unroll
    for (int j = 1; j < 17; j++) tj = (lane == j) ? t[j] : tj;
    int my = 0;
    if (lane < 17) {
        TNode tn;
        make_tnode(tn, ax - tj, s);
        bool inner_ok = true;
        int neg = 0;
        const double val = (kind == 3) ? inner_t<5>(tn, s, sign, &inner_ok, neg) : inner_t<1>(tn, s, sign, &inner_ok, neg);
        my = (inner_ok ? 0 : 1) | ((s.sv_nonzero && neg < 0) ? 2 : 0);
        g[lane] = val * T.inv_wt;
        fl[lane] = my;
    }
    // z-integrals that want more than their first level (integrate.pxi:114-141), one t-node at a time, all lanes on it
    auto redo = [&](unsigned want) {
        unsigned todo = __ballot_sync(0xffffffffu, (lane < 17) && (my & 1) && ((want >> lane) & 1u));
        if (!todo) return;
        double zp[17];
        const NodeTab* Z = T.z[sign];
        dyadic17(zp, Z[0].u[1], Z[1].u[1], Z[2].u[1], Z[3].u[1], Z[4].u[1]);
        double zj = zp[0];
#pragma unroll
        for (int j = 1; j < 17; j++) zj = (lane == j) ? zp[j] : zj;
        while (todo) {
            const int src = __ffs(todo) - 1;
            todo &= todo - 1;
            const double t_src = __shfl_sync(0xffffffffu, tj, src);
            if (T.depth_z == 2) {
                if (lane < 17) gz[lane] = pdf_sv_adhoc(c, ax - t_src, zj) / T.wz[sign];
                __syncwarp();
                unsigned used;
                const double v = walk_depth2(gz, zp, T.wz[sign], T.simps_err, T.depth_z, &used);
                __syncwarp();
                if (lane == src) { g[src] = v * T.inv_wt; fl[src] = 0; my = 0; }
            } else if (lane == src) {   // other budgets: this lane alone, node by node
                const AdhocAxis az{c, ax, tj, true};
                g[src] = adhoc_axis_integral(az, zp[0], zp[16], T.wz[sign], T.simps_err, T.depth_z) * T.inv_wt;
                fl[src] = 0;
                my = 0;
            }
        }
    };
    redo(0x15555u);                    // levels 1 and 2 (even nodes) are visited whenever the first level fails
    __syncwarp();
    unsigned used;
    double result = walk_depth2(g, t, T.wt, T.simps_err, T.depth_t, &used);
    if (__any_sync(0xffffffffu, (lane < 17) && (my & 1) && ((used >> lane) & 1u))) {
        __syncwarp();
        redo(used);                    // third-level nodes of the halves that asked for them
        __syncwarp();
        result = walk_depth2(g, t, T.wt, T.simps_err, T.depth_t, &used);
    }
    int bad = 0;
#pragma unroll
    for (int j = 0; j < 17; j++) bad |= ((used >> j) & 1u) ? fl[j] : 0;
    if (bad & 2) result = CUDART_NAN;
    __syncwarp();   // g / fl are reused for the warp's next trial
    return result;
}

static __device__ __noinline__ double adaptive_refine(double x, const Tables* Tp, int kind, int kl_max)
{
    const Tables& T = *Tp;
    if (kind == 4) return adaptive_full_pdf(x, Tp);   // z-only integration: nothing staged helps below the first level
    const Params& p = T.prm;
    Set s;                                            // private copy: the caller's stays in registers
    rebuild_set(s, Tp, kind, kl_max);
    int neg = 0;
    const double ax = fabs(x);
    const int sign = x > 0;
    if (T.depth_t != 2) return adaptive_full_pdf(x, Tp);   // other budgets: the general walker
    AdhocCtx c;                                           // for z-integrals that need their own refinement
    c.v = sign ? -p.v : p.v; c.v2 = s.v2; c.sv2 = s.sv2; c.a = p.a; c.a2 = s.a2; c.inv_a2 = s.inv_a2;
    c.asv2 = c.a2 * c.sv2; c.sv_nonzero = s.sv_nonzero; c.T = Tp;
    double r = (kind == 3) ? refine_t_depth2<5>(ax, sign, s, c, &neg) : refine_t_depth2<1>(ax, sign, s, c, &neg);
    if (s.sv_nonzero && neg < 0) r = CUDART_NAN;
    return r;
}

// The full reference recursion for one trial: the table-driven walkers for adaptive parameter sets, the
// operation-for-operation evaluator for everything the tables do not cover.
__device__ __forceinline__ double density_generic(double x, const Set& s)
{
    const Tables& T = *s.tab;
    if (s.mode == kModeAdaptive) return adaptive_refine(x, s.tab, s.kind, s.kl_max);
    return generic_full_pdf(x, T.prm, T.o);
}

// Per-trial parameter sets (wiener_like_multi, RL-DDM): nothing can be staged per CTA.  The common regression case
// (sz, st below the 1e-3 threshold) is one node evaluated from registers; everything else takes the generic evaluator.
__device__ __forceinline__ double density_once(double x, const Params& prm, const Opts& o, const KTable& kt)
{
    const double z = prm.z, a = prm.a, t = prm.t, sv = prm.sv, sz = prm.sz, st = prm.st;
    if (kt.n < 0 || kt.kl_max > kMaxKLarge) return generic_full_pdf(x, prm, o);
    // integration over z and / or t with this trial's own parameters: table-driven nodes, the reference's recursion
    // (n == 0 on an integrated axis is the reference's division-by-zero corner, Appendix A.12: operation for operation)
    if (!(sz < 1e-3 && st < 1e-3)) {
        if (!o.use_adaptive && ((sz >= 1e-3 && o.n_sz == 0) || (st >= 1e-3 && o.n_st == 0))) return generic_full_pdf(x, prm, o);
        return full_pdf_adhoc(x, prm, o, &tables());
    }
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
    const int d = klookup_scan(kt, tt);
    const int K = d & 0xff;
    const double c = 0.5 / fma(sv * sv, xp, 1.0);
    const double azsv = a * w * sv;
    const double E = (azsv * azsv - 2 * a * vv * w - prm.v * prm.v * xp) * c;
    const double rs = sqrt(2.0 * c);
    double series, pref;
    if (d >> 8) {
        const double inv_tt = 1.0 / tt;
        series = series_small_slow(w, K, inv_tt, E);
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
