// Lean fixed-grid likelihood kernel: the instantiation the headline benchmark (BASELINE.json configs[3] with configs[1]'s
// fixed 3x3 Simpson grid) and every other fixed-grid launch with n_sz = 2 run through.
//
// Same mathematics, tables and results as like_batch_kernel<FastImpl, false> (wfpt_fast.cuh): the series-term table, the
// staged quadrature nodes, the lattice factorisation of the small-time series, p*exp(E) instead of exp(log p + E).  What
// differs is how the trial loop is organised around the WARP:
//
//  * RTs arrive |rt|-sorted, so the 32 trials a warp holds in one round sit next to each other on the tt axis and nearly
//    always share the series expansion and its term count K.  The warp keeps, per t-node, the piece of the series-term
//    table it was in last time (bounds on the high word of x' = tt a^2, in shared memory); one vote tells whether every
//    lane is still inside.  If so the descriptor is taken through a warp reduction into a uniform register: every branch on
//    "small or large time" and on K is a uniform branch, and each (expansion, K) combination runs straight-line code with
//    exactly its terms.  Otherwise the lanes look their pieces up and the node is evaluated once per distinct descriptor
//    present, lanes of other pieces riding along with their results discarded (two passes at a piece boundary, nothing
//    diverges).
//  * exp arguments are range-checked once per node set through the high words (one three-input max); a node set with an
//    argument outside the branch-free exp kernel's range sends its trial to a short list in shared memory that the CTA
//    evaluates after the loop with the general evaluator (library exp, every term the reference sums) -- no call site and
//    no slow path inside the hot loop.
//  * five resident CTAs per SM at 96 registers; the set's constants live in registers, the tables in 24 KB of shared memory
//    (32 KB in the adaptive first-level kernel), the exp chains of a node set run three or four at a time.
//
// The adaptive launches' two kernels (first level here in the same warp-organised loop, refinement of the parked trials in a
// second kernel) are further down.  DESIGN.md section 4 has the measurements behind each of these choices.
//
// Reference: /root/reference/src/pdf.pxi:28-146 (ftt_01w, pdf_sv, full_pdf), /root/reference/src/integrate.pxi:12-68
// (simpson_1D / simpson_2D), /root/reference/src/wfpt.pyx:54-76 (wiener_like).
#include <cmath>
#include <mutex>

#include "like_kernels.cuh"
// The kernels of this file stage at most five z-nodes per response sign (n_sz == 2 on fixed grids, the five quarter points of
// the first adaptive level): the z tables of fast::Tables shrink accordingly (7.4 KB -> 1.1 KB per CTA; the shared memory not
// taken stays with L1, which holds the RT array and the loop's few spilled values).  Five keeps the two signs' tables
// 560 bytes apart: a warp holding both signs reads them from disjoint banks.
#define WFPT_MAX_ZNODES 5
#include "wfpt_fast.cuh"

namespace wfpt {
namespace lean {

using fast::NodeTab;
using fast::Tables;

constexpr int kWarps = kBlock / 32;
constexpr int kDeferCap = 256;          // trials per CTA the post-loop general evaluator takes before the CTA starts over
constexpr unsigned kFull = 0xffffffffu;
constexpr int kExpBits = 11, kExpN = 1 << kExpBits;   // entries of the exp kernel's 2^(j/N) table

// The set's constants, one copy per thread: the loops read them from registers (reading them from shared memory at their use
// was measured slower).
struct alignas(16) SetConst {
    double a2, inv_a2;        // a^2 and its reciprocal
    double sv2, nv2;          // sv^2, -v^2
    double pref_s, pref_l;    // 1/(sqrt(2 pi) a^2), pi/a^2
    double omp, wp;           // 1 - p_outlier, w_outlier * p_outlier
    double tlo, ct;           // t - st/2 (validity bound on |x|), Simpson scale along t
};

// Cached piece of the series-term table, on the x' axis: a lane whose x' has its high word strictly inside (lo_hi, hi_hi) has
// tt = x'/a^2 (as the reference rounds it) inside the piece -- the bounds are the high words of thr * a^2 moved inwards by
// one unit (2^-20 relative), far more than the rounding of that product and of the quotient can shift anything.
struct alignas(16) PieceCache { int lo_hi, hi_hi, desc, pad; };

struct alignas(16) TRec { double tn, wt; };   // t-node position, and its Simpson weight times the t-axis scale

struct Shared {
    double exptab[kExpN];
    double a2, pad_;               // a^2 for the exact quotient of the piece-boundary passes (read there, not held in a register)
    double pcz[2][2];              // [0][sign]: small-time prefactor 1/(sqrt(2 pi) a^2) times the z-Simpson scale cz[sign]; [1][sign]: pi/a^2 times it
    double gk[2][4];               // per sign: -2 (z_0 + 1), -2 (1 - z_last), -2 step (the exponents of R, Q, rho are these over tt)
    TRec trec[fast::kMaxNodes + 1];
    PieceCache piece[kWarps][fast::kMaxNodes + 1];
    double defer[kDeferCap];
    int xthr_hi[kMaxPieces + 1];   // high words of thr[p] * a^2: the series-term table's thresholds on the x' axis
    int ndefer;
    int nt, sv_nonzero, nzn;
};

// sum_i log(p_i) as log(prod_i p_i) (see LogProduct in like_kernels.cuh), with the zero flag kept inside and a 32-bit exponent
// (a thread sees at most 2^30 trials of at most 2^10 binades each... the slice limit of plan_slices keeps the sum in range).
struct LogProd {
    double m = 1.0, slow = 0.0;
    int e = 0, zero = 0;
    __device__ __forceinline__ void add(double p)
    {
        const int hi = __double2hiint(p);
        if ((unsigned)(hi - 0x00200000) < 0x7fb00000u) {  // 2^-1020 <= p < 2^1022
            m *= p;
            const int hm = __double2hiint(m);
            e += (hm >> 20) - 1023;
            m = __hiloint2double((hm & 0x000fffff) | 0x3ff00000, __double2loint(m));
        } else {
            add_slow(p);
        }
    }
    __device__ __forceinline__ void add_slow(double p)
    {
        zero |= (p == 0);
        slow += log(p);
    }
    __device__ __forceinline__ double value() const { return slow + fma((double)e, 0.6931471805599453, log(m)); }
};

__device__ __forceinline__ Shared& shared()
{
    __shared__ Shared s;
    return s;
}

// A value that is the same in every lane, taken through a warp reduction: it lands in a uniform register, and branches and
// loop counters that hang on it become uniform ones (no reconvergence bookkeeping, compares on the uniform datapath).  All
// 32 lanes must be there.
__device__ __forceinline__ int uni(int v) { return __reduce_max_sync(kFull, v); }

// The exp table travels global -> shared by cp.async, issued before the set's tables are computed and awaited behind them:
// its L2 latency is off the CTA's prologue (a prologue's duration is time in which the CTA contributes no warps).
__device__ __forceinline__ void exptab_fetch(double* dst, const double* src)
{
    const unsigned d0 = (unsigned)__cvta_generic_to_shared(dst);
    for (int i = threadIdx.x; i < kExpN / 2; i += kBlock)
        asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(d0 + 16u * i), "l"(src + 2 * i) : "memory");
}
__device__ __forceinline__ void exptab_wait() { asm volatile("cp.async.wait_all;" ::: "memory"); }

// ---- shared memory by 32-bit address ------------------------------------------------------------------------------------------
// The tables are read through explicit shared-window addresses (one base per CTA, member offsets as immediates): left to
// itself the compiler re-derives the window base (cluster CTA id, shift, add) at most use sites, a dozen uniform-datapath
// instructions per t-node that all take issue slots.  The loads are declared without side effects: the tables do not
// change once the CTA has passed the barrier behind prepare().
template <int OFF>
__device__ __forceinline__ double2 lds2(unsigned a)
{
    double2 v;
    asm("ld.shared.v2.f64 {%0, %1}, [%2+%3];" : "=d"(v.x), "=d"(v.y) : "r"(a), "n"(OFF));
    return v;
}
template <int OFF>
__device__ __forceinline__ double lds1(unsigned a)
{
    double v;
    asm("ld.shared.f64 %0, [%1+%2];" : "=d"(v) : "r"(a), "n"(OFF));
    return v;
}
// volatile form: for a value that must be read where it is used (a loop-invariant one would otherwise be hoisted into
// registers for the whole loop)
template <int OFF>
__device__ __forceinline__ double lds1v(unsigned a)
{
    double v;
    asm volatile("ld.shared.f64 %0, [%1+%2];" : "=d"(v) : "r"(a), "n"(OFF));
    return v;
}
// a^2's offset from the exp table's address (the one shared-memory base the node routines are handed)
constexpr int kScA2 = (int)offsetof(Shared, a2) - (int)offsetof(Shared, exptab);
constexpr int kPczS = (int)offsetof(Shared, pcz) - (int)offsetof(Shared, exptab), kPczL = kPczS + 16,
              kGk = (int)offsetof(Shared, gk) - (int)offsetof(Shared, exptab);
constexpr int kOffZ = (int)offsetof(Tables, z);
constexpr int kNodeBytes = (int)sizeof(NodeTab), kSignBytes = kNodeBytes * fast::kMaxZNodes;
constexpr int kNA = (int)offsetof(NodeTab, A), kNU = (int)offsetof(NodeTab, u), kNS = (int)offsetof(NodeTab, S);
static_assert(kNA % 16 == 0 && kNU % 16 == 0 && kNS % 16 == 0 && kNodeBytes % 16 == 0, "NodeTab pairs must be 16-byte aligned");

// ---- exp, three or four independent chains side by side -------------------------------------------------------------------
// exp(x) = 2^e * 2^(j/2048) * exp(r),  k = rint(2048 x / ln 2) = 2048 e + j,  r = x - k ln2/2048, |r| <= 1.7e-4, so that
// exp(r) - 1 = r + r^2 (1/2 + r/6) to 3.4e-17: seven FP64 instructions per exp, six deep (the 64-entry form of
// wfpt_fast.cuh needs ten).  The table (16 KB per CTA, correctly rounded entries computed once on the host) is the price.
// The reduction multiplies k by ln2/2048 rounded to double in ONE fused step: its rounding (2^-54 relative) enters as
// exp(x (1 + 2^-54)), i.e. as half a unit in the last place of the ARGUMENT -- which every argument here already carries
// from the operations that formed it; a two-constant reduction would buy nothing but one more link in the chain
// (measured: 18.7 -> 17.5 ms on the headline launch, the reported parity errors unchanged to all printed digits).  |x| < 708.
__constant__ double kE[5] = {2954.639443740597, 6755399441055744.0, -0.6931471805599453094172321 / 2048.0, 1.0 / 6.0, 0.5};
__device__ __align__(16) double g_exp_table[kExpN];   // 2^(j/2048) less j * 2^9 in the high word (see expv); filled by ensure_exp_table()

template <int M>
__device__ __forceinline__ void expv(const double (&x)[M], double (&out)[M], unsigned tab)
{
    double t[M], r[M];
    int ki[M];
#pragma unroll
    for (int i = 0; i < M; i++) t[i] = fma(x[i], kE[0], kE[1]);
#pragma unroll
    for (int i = 0; i < M; i++) {
        ki[i] = __double2loint(t[i]);
        t[i] = t[i] - kE[1];
    }
#pragma unroll
    for (int i = 0; i < M; i++) r[i] = fma(t[i], kE[2], x[i]);
#pragma unroll
    for (int i = 0; i < M; i++) t[i] = fma(r[i], kE[3], kE[4]);
#pragma unroll
    for (int i = 0; i < M; i++) t[i] = fma(t[i], r[i] * r[i], r[i]);
#pragma unroll
    for (int i = 0; i < M; i++) {
        // table entry j holds 2^(j/2048) with j * 2^9 taken off its high word, so that adding k * 2^9 = (e << 20) + j * 2^9
        // yields S = 2^e * 2^(j/2048) in one integer multiply-add, off the chain that ends in the product below
        const double Tb = lds1<0>(tab + ((ki[i] & (kExpN - 1)) << 3));
        int hi;
        asm("mad.lo.s32 %0, %1, 512, %2;" : "=r"(hi) : "r"(ki[i]), "r"(__double2hiint(Tb)));
        const double S = __hiloint2double(hi, __double2loint(Tb));
        out[i] = fma(S, t[i], S);
    }
}

// non-positive argument whose exp may underflow: anything below -707 is taken as -707 (the result only ever multiplies
// terms that are then 1e-307 of the leading one)
__device__ __forceinline__ double clamp_neg(double x)
{
    const unsigned hi = min((unsigned)__double2hiint(x), 0xC0861800u);
    return __hiloint2double((int)hi, __double2loint(x));
}
__device__ __forceinline__ int abs_hi(double x) { return __double2hiint(x) & 0x7fffffff; }
constexpr int kExpHiLimit = 0x40862000;   // |x| < 708 <=> abs_hi(x) < this (NaN and inf fail too)

// 1/sqrt(x), normal positive x: hardware seed (about 2^-23 accurate), one third-order step (-> 2^-66), five FP64 instructions
__device__ __forceinline__ double rsqrt_d(double x)
{
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    const double e = fma(-x * y, y, 1.0);                 // 1 - x y^2
    const double p = fma(e, 0.375, 0.5) * e;              // e/2 + 3 e^2/8
    return fma(y, p, y);
}

// ---- one t-node: inner z-integral (or the single node) for a warp-uniform descriptor ---------------------------------------
// d = (small << 8) | K, uniform over the warp.  Lanes whose own descriptor differs, or that carry no live node (x' <= 0),
// compute along and are ignored by the caller.  Returns the Simpson sum over z times cz (NZN == 3) or the node (NZN == 1);
// ored into `neg`: sign bits of the series sums; `bad` |= an exp argument outside the fast kernel's range.
template <int NZN, bool SV>
__device__ __forceinline__ double tnode_value(int d, double xp, double tt, double rs, double c, unsigned tb, unsigned zb,
                                              unsigned tab, const SetConst& sc, int sign, int& neg, int& bad)
{
    // rs = 1/sqrt(sv^2 x' + 1), c = rs^2 / 2 (1 and 1/2 without drift variability); tt = x'/a^2 to the last bit or so (the
    // series-term decision was taken by the caller on exact grounds)
    const double E0 = sc.nv2 * xp * c;
    double y[NZN];
    const int K = d & 0xff;
    double pref;
    if (d >> 8) {
        const double r = rsqrt_d(tt);
        const double tau = r * r, m2t = -2.0 * tau;
        pref = (NZN == 3) ? tau * r * rs * lds1<kPczS>(tab + 8 * sign) : tau * r * sc.pref_s * rs;
        double a[NZN], e0[NZN];
        {
            const double2 Au0 = lds2<kNA>(zb);                                // (A, -z^2/2) of node 0, 1, 2
            a[0] = fma(Au0.y, tau, fma(Au0.x, c, E0));
            if (NZN == 3) {
                const double2 Au1 = lds2<kNA + kNodeBytes>(zb), Au2 = lds2<kNA + 2 * kNodeBytes>(zb);
                a[1] = fma(Au1.y, tau, fma(Au1.x, c, E0));
                a[NZN - 1] = fma(Au2.y, tau, fma(Au2.x, c, E0));
            }
        }
        if (NZN == 3) bad |= (max(max(abs_hi(a[0]), abs_hi(a[1])), abs_hi(a[NZN - 1])) >= kExpHiLimit);
        else bad |= (abs_hi(a[0]) >= kExpHiLimit);
        expv<NZN>(a, e0, tab);
        if (NZN == 3) {
            const double2 ua0 = lds2<kNU>(zb), ua1 = lds2<kNU + kNodeBytes>(zb), ua2 = lds2<kNU + 2 * kNodeBytes>(zb);   // (z-2, z)
            const double2 ub0 = lds2<kNU + 16>(zb), ub1 = lds2<kNU + 16 + kNodeBytes>(zb), ub2 = lds2<kNU + 16 + 2 * kNodeBytes>(zb);   // (z+2, z+4)
            const double um2[3] = {ua0.x, ua1.x, ua2.x}, u0[3] = {ua0.y, ua1.y, ua2.y}, up2[3] = {ub0.x, ub1.x, ub2.x},
                         up4[3] = {ub0.y, ub1.y, ub2.y};
            double g[3], w[3];
            {
                const double2 k01 = lds2<kGk>(tab + 32 * sign);
                g[0] = clamp_neg(tau * k01.x);                                  // R   = exp(-2 (z_0 + 1) / tt)
                g[1] = clamp_neg(tau * k01.y);                                  // Q   = exp(-2 (1 - z_last) / tt)
                g[2] = clamp_neg(tau * lds1<kGk + 16>(tab + 32 * sign));        // rho = exp(-2 step / tt)
            }
            expv<3>(g, w, tab);
            const double rho = w[2];
            double Rn[3], Qn[3];
            Rn[0] = w[0];
            Rn[1] = Rn[0] * rho;
            Rn[2] = Rn[1] * rho;
            Qn[2] = (K >= 3) ? w[1] : 0.0;                     // K == 2 sums k = 0, 1 only
            Qn[1] = Qn[2] * rho;
            Qn[0] = Qn[1] * rho;
#pragma unroll
            for (int n = 0; n < 3; n++) y[n] = fma(um2[n], Qn[n], fma(up2[n], Rn[n], u0[n]));
            if (K >= 4) {                                      // uniform: k = +2 term, e(k=2)/e(k=1) = R_n * exp(-4/tt)
                const double H = w[0] * (w[1] * rho * rho);
#pragma unroll
                for (int n = 0; n < 3; n++) y[n] = fma(up4[n], Rn[n] * Rn[n] * H, y[n]);
            }
#pragma unroll
            for (int n = 0; n < 3; n++) y[n] = fma(e0[n], y[n], 0.0);
        } else {
            const double2 ua = lds2<kNU>(zb), ub = lds2<kNU + 16>(zb);
            double g[2], w[2];
            g[0] = clamp_neg(fma(m2t, ua.y, m2t));
            g[1] = clamp_neg(fma(-m2t, ua.y, m2t));
            expv<2>(g, w, tab);
            const double R = w[0], Q = (K >= 3) ? w[1] : 0.0;
            double inner = fma(ua.x, Q, fma(ub.x, R, ua.y));
            if (K >= 4) inner = fma(ub.y, R * R * (w[0] * w[1]), inner);
            y[0] = fma(e0[0], inner, 0.0);
        }
    } else {
        pref = (NZN == 3) ? rs * lds1<kPczL>(tab + 8 * sign) : sc.pref_l * rs;
        double a[NZN + 1], w[NZN + 1];
        a[NZN] = -fast::kHalfPiSq * tt;
        a[0] = fma(lds1<kNA>(zb), c, E0);
        if (NZN == 3) {
            a[1] = fma(lds1<kNA + kNodeBytes>(zb), c, E0);
            a[NZN - 1] = fma(lds1<kNA + 2 * kNodeBytes>(zb), c, E0);
        }
        if (NZN == 3) bad |= (max(max(max(abs_hi(a[0]), abs_hi(a[1])), abs_hi(a[2])), abs_hi(a[3])) >= kExpHiLimit);
        else bad |= (max(abs_hi(a[0]), abs_hi(a[1])) >= kExpHiLimit);
        expv<NZN + 1>(a, w, tab);
        const double q = w[NZN], q2 = q * q, q4 = q2 * q2;
        {
            const double2 s0 = lds2<kNS>(zb);                                  // (sin, 2 sin 2) of node 0, 1, 2
            y[0] = q * s0.x;
            if (K >= 2) y[0] = fma(q4, s0.y, y[0]);
            if (NZN == 3) {
                const double2 s1 = lds2<kNS + kNodeBytes>(zb), s2 = lds2<kNS + 2 * kNodeBytes>(zb);
                y[1] = q * s1.x;
                y[NZN - 1] = q * s2.x;
                if (K >= 2) {
                    y[1] = fma(q4, s1.y, y[1]);
                    y[NZN - 1] = fma(q4, s2.y, y[NZN - 1]);
                }
            }
        }
        if (K >= 3) {
            const double q8 = q4 * q4, q9 = q8 * q, q16 = (K >= 4) ? q8 * q8 : 0.0;
            const double2 s0 = lds2<kNS + 16>(zb);                             // (3 sin 3, 4 sin 4)
            y[0] = fma(q16, s0.y, fma(q9, s0.x, y[0]));
            if (NZN == 3) {
                const double2 s1 = lds2<kNS + 16 + kNodeBytes>(zb), s2 = lds2<kNS + 16 + 2 * kNodeBytes>(zb);
                y[1] = fma(q16, s1.y, fma(q9, s1.x, y[1]));
                y[NZN - 1] = fma(q16, s2.y, fma(q9, s2.x, y[NZN - 1]));
            }
        }
#pragma unroll
        for (int n = 0; n < NZN; n++) y[n] = fma(y[n], w[n], 0.0);
    }
#pragma unroll
    for (int n = 0; n < NZN; n++) neg |= __double2hiint(y[n]);   // the prefactor is positive: the sign is the series sum's
    // integrate.pxi:30-45 with n_sz = 2; prefactor and Simpson scale hz/(3 wz) applied to the sum as one multiplier
    if (NZN == 3) return (fma(4.0, y[1], y[0]) + y[NZN - 1]) * pref;
    return y[0] * pref;
}

// ---- one t-node of the first ADAPTIVE level: five equally spaced z-nodes ------------------------------------------------------
// integrate.pxi:72-141 with the five quarter points of [z - sz/2, z + sz/2] staged.  The exponent of node n is quadratic in
// n (a_n = a_0 + n beta + n^2 gamma), so e0_{n+1}/e0_n = exp(beta + gamma) exp(2 gamma)^n: three exps for the five leading
// terms (small time) or the five exp(E_n) (large time).  Returns the first-level value of the z-integral; *deeper |= the
// reference's recursion would go below the first level here (the trial is then parked for the refinement routines).
constexpr int kOffZq = (int)offsetof(Tables, zq), kOffWz = (int)offsetof(Tables, wz), kOffIwz = (int)offsetof(Tables, inv_wz);

template <bool SV>
__device__ __forceinline__ double tnode_value5(int d, double xp, double tt, double rs, double c, unsigned tb, unsigned zb,
                                               unsigned tab, const SetConst& sc, int sign, double simps_err, int depth_z,
                                               int& neg, int& bad, bool& deeper)
{
    const double E0 = sc.nv2 * xp * c;
    const int K = d & 0xff;
    // (the quadratic-exponent coefficients are 8-byte aligned in fast::Tables: single loads)
    const double2 zq01 = make_double2(lds1<kOffZq>(tb + 32 * sign), lds1<kOffZq + 8>(tb + 32 * sign));
    const double2 zq23 = make_double2(lds1<kOffZq + 16>(tb + 32 * sign), lds1<kOffZq + 24>(tb + 32 * sign));
    double y[5];
    double pref;
    if (d >> 8) {
        const double r = rsqrt_d(tt);
        const double tau = r * r;
        pref = tau * r * sc.pref_s * rs;
        const double gam = fma(zq01.y, c, zq01.x * tau), bet = fma(zq23.y, c, zq23.x * tau);
        const double2 Au0 = lds2<kNA>(zb);
        const double2 ua0 = lds2<kNU>(zb), ua4 = lds2<kNU + 4 * kNodeBytes>(zb);          // (z - 2, z) of the end nodes
        // two batches of three exps, node values streamed: the leading terms run up the nodes (e0, R_n by running products),
        // the k = -1 terms run down them (Q_n); only y[] and e0_n (z_n - 2) stay live in between
        double g[3], w[3];
        g[0] = fma(Au0.y, tau, fma(Au0.x, c, E0));
        g[1] = bet + gam;                                               // a_1 - a_0
        g[2] = 2.0 * gam;
        const double a4 = fma(4.0, bet, fma(16.0, gam, g[0]));           // exponent of the last node
        // every e0_n and every ratio between neighbours must be inside the exp kernel's range
        bad |= (max(abs_hi(g[0]), abs_hi(a4)) >= kExpHiLimit) | !(fabs(g[1]) < 700.0) | !(fabs(fma(7.0, gam, bet)) < 700.0);
        double h[3], u[3];
        {
            const double2 k01 = lds2<kGk>(tab + 32 * sign);
            h[0] = clamp_neg(tau * k01.x);                                  // R   = exp(-2 (z_0 + 1) / tt)
            h[1] = clamp_neg(tau * k01.y);                                  // Q   = exp(-2 (1 - z_last) / tt)
            h[2] = clamp_neg(tau * lds1<kGk + 16>(tab + 32 * sign));        // rho = exp(-2 step / tt)
        }
        expv<3>(g, w, tab);
        expv<3>(h, u, tab);
        const double rho = u[2], rho2 = rho * rho;
        const double H = u[0] * (u[1] * (rho2 * rho2));                  // R_0 Q_0 = exp(-4/tt), for the k = +2 term
        double v[5];
        {
            double e = w[0], ratio = w[1], Rn = u[0], R2H = u[0] * u[0] * H;     // R_n^2 H runs up the nodes by rho^2
#pragma unroll
            for (int n = 0; n < 5; n++) {
                const double2 ua = (n == 0) ? ua0 : (n == 4) ? ua4 : lds2<kNU>(zb + n * kNodeBytes);
                const double2 ub = lds2<kNU + 16>(zb + n * kNodeBytes);  // (z + 2, z + 4)
                double t = fma(ub.x, Rn, ua.y);
                if (K >= 4) {
                    t = fma(ub.y, R2H, t);
                    R2H *= rho2;
                }
                y[n] = fma(e, t, 0.0);                                  // (+ 0.0: an underflowed e0 times a negative sum gives
                v[n] = e * ua.x;                                        //  +0, not -0, and -0 * Q + (+0) below stays +0)
                e *= ratio;
                ratio *= w[2];
                Rn *= rho;
            }
        }
        if (K >= 3) {                                                    // (K == 2 sums k = 0, 1 only)
            double Qn = u[1];
#pragma unroll
            for (int n = 4; n >= 0; n--) {
                y[n] = fma(v[n], Qn, y[n]);
                Qn *= rho;
            }
        }
    } else {
        pref = sc.pref_l * rs;
        const double gam = zq01.y * c, bet = zq23.y * c;
        double g[4], w[4];
        g[0] = fma(lds1<kNA>(zb), c, E0);
        g[1] = bet + gam;
        g[2] = 2.0 * gam;
        g[3] = -fast::kHalfPiSq * tt;
        const double a4 = fma(4.0, bet, fma(16.0, gam, g[0]));
        bad |= (max(max(abs_hi(g[0]), abs_hi(a4)), abs_hi(g[3])) >= kExpHiLimit) | !(fabs(g[1]) < 700.0) |
               !(fabs(fma(7.0, gam, bet)) < 700.0);
        expv<4>(g, w, tab);
        const double q = w[3], q2 = q * q, q4 = q2 * q2, q8 = q4 * q4;
        const double q9 = (K >= 3) ? q8 * q : 0.0, q16 = (K >= 4) ? q8 * q8 : 0.0, q4k = (K >= 2) ? q4 : 0.0;
        double ex = w[0], ratio = w[1];
#pragma unroll
        for (int n = 0; n < 5; n++) {
            const double2 s01 = lds2<kNS>(zb + n * kNodeBytes);
            double acc = fma(q4k, s01.y, q * s01.x);
            if (K >= 3) {
                const double2 s23 = lds2<kNS + 16>(zb + n * kNodeBytes);
                acc = fma(q16, s23.y, fma(q9, s23.x, acc));
            }
            y[n] = fma(acc, ex, 0.0);
            ex *= ratio;
            ratio *= w[2];
        }
    }
#pragma unroll
    for (int n = 0; n < 5; n++) neg |= __double2hiint(y[n]);   // the prefactor is positive: the sign is the series sum's
    // integrate.pxi:99-105 (first level) on the five node values; prefactor, 1/width and h/12 go onto the three sums as one
    // multiplier instead of onto the five values
    const double s12 = pref * (lds1<kOffIwz>(tb + 8 * sign) * (lds1<kOffWz>(tb + 8 * sign) * (1.0 / 12.0)));
    const double S = (2.0 * s12) * (y[0] + 4 * y[2] + y[4]);
    const double Sl = s12 * (y[0] + 4 * y[1] + y[2]);
    const double Sr = s12 * (y[2] + 4 * y[3] + y[4]);
    const double S2 = Sl + Sr;
    deeper |= !((depth_z <= 0) || (fabs(S2 - S) <= 15 * simps_err));
    return fma(S2 - S, 1.0 / 15.0, S2);
}

// Per-lane table lookup (slow path only).
__device__ __forceinline__ int lookup_lane(const Tables& T, double tt, int* piece)
{
    const long long b = __double_as_longlong(tt);
    int p = 0;
    const int n = T.npieces;
    while (p < n && b >= T.thr[p]) p++;
    *piece = p;
    return T.desc[p];
}

// Density of the warp's 32 trials (one per lane; `live` lanes only) on the staged fixed grid.
template <int NZN, bool SV>
__device__ __forceinline__ double warp_density(double ax, int sign, bool live, const Tables& T, Shared& S, const SetConst& sc,
                                               int nt, unsigned tb, unsigned etab, int& neg, int& bad)
{
    const int w = threadIdx.x >> 5;
    const unsigned zb = tb + kOffZ + sign * kSignBytes;
    double acc = 0.0;
    // the record of the next t-node (position, weight, cached table piece) is fetched one node ahead: shared-memory latency
    // stays off the dependent chain
    TRec rec_n = S.trec[0];
    PieceCache pc_n = S.piece[w][0];
#pragma unroll 1
    for (int i = 0; i <= nt; i++) {
        const TRec rec = rec_n;
        const PieceCache pc = pc_n;
        rec_n = S.trec[i + 1];
        pc_n = S.piece[w][i + 1];
        const double xp = ax - rec.tn;
        const bool alive = live && (xp > 0);
        const int th = __double2hiint(xp);
        double val = 0.0;
        const unsigned am = __ballot_sync(kFull, alive);
        if (am == 0) continue;                                  // nobody has this node (x' <= 0 everywhere): contributes 0
        const double tt = xp * sc.inv_a2;
        double rs = 1.0, c = 0.5;
        if (SV) {
            rs = rsqrt_d(fma(sc.sv2, xp, 1.0));
            c = 0.5 * rs * rs;
        }
        if (__all_sync(kFull, !alive || (th > pc.lo_hi && th < pc.hi_hi))) {
            int ng = 0, bd = 0;
            val = tnode_value<NZN, SV>(uni(pc.desc), xp, tt, rs, c, tb, zb, etab, sc, sign, ng, bd);
            if (alive) { neg |= ng; bad |= bd; }                // lanes without the node computed along on garbage
        } else {
            // a piece boundary runs through the warp (or the warp has moved on): look up per lane on the reference's own
            // tt = x'/pow(a, 2) (pdf.pxi:81,98), evaluate once per distinct descriptor present, and remember the piece of the
            // last live lane for the next round
            int piece = 0;
            const int mine = alive ? lookup_lane(T, __ddiv_rn(xp, lds1v<kScA2>(etab)), &piece) : -1;
            unsigned todo = am;
            int neg_l = 0, bad_l = 0;
            while (todo) {
                const int src = __ffs(todo) - 1;
                const int d = __shfl_sync(kFull, mine, src);
                const bool match = (mine == d);
                int ng = 0, bd = 0;
                const double v = tnode_value<NZN, SV>(d, xp, tt, rs, c, tb, zb, etab, sc, sign, ng, bd);
                if (match) { val = v; neg_l = ng; bad_l = bd; }
                todo &= ~__ballot_sync(kFull, match);
            }
            neg |= neg_l;
            bad |= bad_l;
            {
                const int last = 31 - __clz(am);
                const int p = __shfl_sync(kFull, piece, last);
                if ((threadIdx.x & 31) == 0) {
                    PieceCache nc;
                    nc.lo_hi = (p > 0) ? S.xthr_hi[p - 1] + 1 : (int)0x80000000;
                    nc.hi_hi = (p < T.npieces) ? S.xthr_hi[p] - 1 : 0x7fffffff;
                    nc.desc = T.desc[p];
                    nc.pad = 0;
                    S.piece[w][i] = nc;
                }
                __syncwarp();
            }
        }
        if (!alive) val = 0.0;
        acc = fma(rec.wt, val, acc);
    }
    return acc;
}

// The common grid, n_st = 2: three t-nodes.  Everything the three nodes need before their series -- x', the votes on
// the cached table pieces, tt, 1/sqrt(sv^2 x' + 1) -- is formed for all three side by side (three independent chains instead
// of one after the other), then the three node sets run back to back.  Returns false (nothing computed) unless the whole
// warp is in the plain case; the caller then takes the node-by-node routine above, which masks and looks pieces up.
template <int NZN, bool SV>
__device__ __forceinline__ bool warp_density3(double ax, int sign, bool live, Shared& S, const SetConst& sc, unsigned tb,
                                              unsigned etab, double* out, int& neg, int& bad)
{
    const int w = threadIdx.x >> 5;
    const unsigned zb = tb + kOffZ + sign * kSignBytes;
    TRec rec[3];
    PieceCache pc[3];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        rec[i] = S.trec[i];
        pc[i] = S.piece[w][i];
    }
    // only the plain case runs here: every lane carries a trial with all three nodes alive, each inside its cached piece --
    // no masks, no per-node votes.  (Anything else: the node-by-node routine.)
    double xp[3];
    bool good = live;
#pragma unroll
    for (int i = 0; i < 3; i++) {
        xp[i] = ax - rec[i].tn;
        const int th = __double2hiint(xp[i]);
        good &= (xp[i] > 0) && (th > pc[i].lo_hi && th < pc[i].hi_hi);
    }
    if (!__all_sync(kFull, good)) return false;
    double tt[3], rs[3], c[3];
#pragma unroll
    for (int i = 0; i < 3; i++) {
        tt[i] = xp[i] * sc.inv_a2;
        rs[i] = 1.0;
        c[i] = 0.5;
    }
    if (SV) {
        double u[3], y[3];
#pragma unroll
        for (int i = 0; i < 3; i++) u[i] = fma(sc.sv2, xp[i], 1.0);
#pragma unroll
        for (int i = 0; i < 3; i++) asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y[i]) : "d"(u[i]));
#pragma unroll
        for (int i = 0; i < 3; i++) u[i] = fma(-u[i] * y[i], y[i], 1.0);
#pragma unroll
        for (int i = 0; i < 3; i++) rs[i] = fma(y[i], fma(u[i], 0.375, 0.5) * u[i], y[i]);
#pragma unroll
        for (int i = 0; i < 3; i++) c[i] = 0.5 * rs[i] * rs[i];
    }
    double acc = 0.0;
    // the descriptors are the same in every lane (one cached piece per warp and node): as uniform values every branch on
    // the expansion and on K is a uniform branch
    int du[3];
    {
        const int packed = uni(pc[0].desc | (pc[1].desc << 10) | (pc[2].desc << 20));   // (a descriptor has nine bits)
        du[0] = packed & 0x3ff; du[1] = (packed >> 10) & 0x3ff; du[2] = packed >> 20;
    }
#pragma unroll
    for (int i = 0; i < 3; i++)
        acc = fma(rec[i].wt, tnode_value<NZN, SV>(du[i], xp[i], tt[i], rs[i], c[i], tb, zb, etab, sc, sign, neg, bad), acc);
    *out = acc;
    return true;
}

// Out-of-line copy of the node evaluation for the piece-boundary passes (cold on sorted inputs): keeps the loop's hot path the
// only inlined copy.  flags: bit 31.. = sign bits seen (neg), bit 0 = bad, bit 1 = deeper.
struct NodeOut { double val; int neg; int flags; };
template <int NZN, bool SV>
static __device__ __noinline__ NodeOut node_call(int d, double xp, double tt, double rs, double c, unsigned tb, unsigned zb,
                                                 unsigned etab, double a2, double inv_a2, double sv2, double nv2, double pref_s,
                                                 double pref_l, int sign, double simps_err, int depth_z)
{
    SetConst sc;
    sc.a2 = a2; sc.inv_a2 = inv_a2; sc.sv2 = sv2; sc.nv2 = nv2; sc.pref_s = pref_s; sc.pref_l = pref_l;
    sc.omp = sc.wp = sc.tlo = sc.ct = 0.0;
    NodeOut o;
    int ng = 0, bd = 0;
    bool dp = false;
    if (NZN == 5) o.val = tnode_value5<SV>(d, xp, tt, rs, c, tb, zb, etab, sc, sign, simps_err, depth_z, ng, bd, dp);
    else o.val = tnode_value<1, SV>(d, xp, tt, rs, c, tb, zb, etab, sc, sign, ng, bd);
    o.neg = ng;
    o.flags = (bd ? 1 : 0) | (dp ? 2 : 0);
    return o;
}

// First adaptive level of the warp's 32 trials (integrate.pxi:72-206): five t-nodes (or the single one when the t axis is not
// integrated), each with the five-node z-integral (or the single node).  *deeper |= the reference would refine this trial
// below the first level, along t or inside some z-integral: the caller parks it.  The three Simpson sums of the t level are
// accumulated node by node in the reference's order of additions.
constexpr int kOffWtA = (int)offsetof(Tables, wt), kOffIwt = (int)offsetof(Tables, inv_wt);

template <int NZN, bool SV>
__device__ __forceinline__ double warp_density_adapt(double ax, int sign, bool live, const Tables& T, Shared& S,
                                                     const SetConst& sc, int nt, unsigned tb, unsigned etab, double simps_err,
                                                     int depth_z, int depth_t, int& neg, int& bad, bool& deeper)
{
    const int w = threadIdx.x >> 5;
    const unsigned zb = tb + kOffZ + sign * kSignBytes;
    const double inv_wt = (nt > 0) ? lds1<kOffIwt>(tb) : 1.0;
    double Sa = 0.0, Sl = 0.0, Sr = 0.0, single = 0.0;
    TRec rec_n = S.trec[0];
    PieceCache pc_n = S.piece[w][0];
#pragma unroll 1
    for (int i = 0; i <= nt; i++) {
        const TRec rec = rec_n;
        const PieceCache pc = pc_n;
        rec_n = S.trec[i + 1];
        pc_n = S.piece[w][i + 1];
        const double xp = ax - rec.tn;
        const bool alive = live && (xp > 0);
        const int th = __double2hiint(xp);
        double val = 0.0;
        const unsigned am = __ballot_sync(kFull, alive);
        if (am != 0) {
            const double tt = xp * sc.inv_a2;
            double rs = 1.0, c = 0.5;
            if (SV) {
                rs = rsqrt_d(fma(sc.sv2, xp, 1.0));
                c = 0.5 * rs * rs;
            }
            auto node = [&](int d, int& ng1, int& bd1, bool& dp1) -> double {
                if (NZN == 5) return tnode_value5<SV>(d, xp, tt, rs, c, tb, zb, etab, sc, sign, simps_err, depth_z, ng1, bd1, dp1);
                return tnode_value<1, SV>(d, xp, tt, rs, c, tb, zb, etab, sc, sign, ng1, bd1);
            };
            if (__all_sync(kFull, !alive || (th > pc.lo_hi && th < pc.hi_hi))) {
                int ng = 0, bd = 0;
                bool dp = false;
                val = node(pc.desc, ng, bd, dp);
                if (alive) { neg |= ng; bad |= bd; deeper |= dp; }    // lanes without the node computed along on garbage
                else val = 0.0;
            } else {
                // a piece boundary runs through the warp (or the warp has moved on): per-lane lookups on the reference's own
                // tt = x'/pow(a, 2), one pass per distinct descriptor, lanes of other pieces riding along
                int piece = 0;
                const int mine = alive ? lookup_lane(T, __ddiv_rn(xp, lds1v<kScA2>(etab)), &piece) : -1;
                unsigned todo = am;
#pragma unroll 1
                while (todo) {
                    const int d = __shfl_sync(kFull, mine, __ffs(todo) - 1);
                    const NodeOut o = node_call<NZN, SV>(d, xp, tt, rs, c, tb, zb, etab, sc.a2, sc.inv_a2, sc.sv2, sc.nv2, sc.pref_s,
                                                         sc.pref_l, sign, simps_err, depth_z);
                    const bool match = (mine == d);
                    if (match) { val = o.val; neg |= o.neg; bad |= o.flags & 1; deeper |= (o.flags & 2) != 0; }
                    todo &= ~__ballot_sync(kFull, match);
                }
                const int p = __shfl_sync(kFull, piece, 31 - __clz(am));   // remember the piece of the last live lane
                if ((threadIdx.x & 31) == 0) {
                    PieceCache nc;
                    nc.lo_hi = (p > 0) ? S.xthr_hi[p - 1] + 1 : (int)0x80000000;
                    nc.hi_hi = (p < T.npieces) ? S.xthr_hi[p] - 1 : 0x7fffffff;
                    nc.desc = T.desc[p];
                    nc.pad = 0;
                    S.piece[w][i] = nc;
                }
                __syncwarp();
            }
        }
        // integrate.pxi:99-105 (first level): S = f0 + 4 f2 + f4, Sleft = f0 + 4 f1 + f2, Sright = f2 + 4 f3 + f4
        const double g = val;   // (1 / width of the t interval goes onto the sums below)
        if (i == 0) { Sa = g; Sl = g; single = val; }
        else if (i == 1) Sl = fma(4.0, g, Sl);
        else if (i == 2) { Sa = fma(4.0, g, Sa); Sl += g; Sr = g; }
        else if (i == 3) Sr = fma(4.0, g, Sr);
        else { Sa += g; Sr += g; }
    }
    if (nt == 0) return single;
    const double h12 = inv_wt * (lds1<kOffWtA>(tb) * (1.0 / 12.0));
    const double S1 = (2.0 * h12) * Sa, S2 = h12 * Sl + h12 * Sr;
    deeper |= !((depth_t <= 0) || (fabs(S2 - S1) <= 15 * simps_err));
    return fma(S2 - S1, 1.0 / 15.0, S2);
}

__device__ __forceinline__ bool lean_set_ok(int mode, int kind)
{
    return mode == fast::kModeZero || mode == fast::kModePoint || (mode == fast::kModeFixed && (kind == 0 || kind == 1));
}

// General evaluator for the handful of trials the loop defers, and for the whole slice when the list overflows.
static __device__ __noinline__ double general_density(double x, int kind, int nz, int nt, int kl_max, int mode, double tlo)
{
    fast::Set s;
    fast::rebuild_set(s, &fast::tables(), kind, kl_max);
    s.mode = mode;
    s.nz = nz;
    s.nt = nt;
    s.tlo = tlo;
    bool defer;
    return fast::density<false>(x, s, &defer);
}

__global__ void __launch_bounds__(kBlock, 5) like_lean_kernel(const double* __restrict__ x, int64_t n,
                                                             const double* __restrict__ params, int64_t P, Opts o,
                                                             double w_outlier, int slices, int64_t chunk,
                                                             double* __restrict__ out_sum, int* __restrict__ out_zero,
                                                             int* __restrict__ done, double* __restrict__ final_out,
                                                             const __grid_constant__ KTable kt)
{
    const int64_t bid = blockIdx.x;
    const int64_t p = bid / slices;
    const int s = (int)(bid - p * slices);
    if (p >= P) return;
    const double* row = params + 8 * p;
    Params prm{row[0], row[1], row[2], row[3], row[4], row[5], row[6], row[7]};
    const double po = prm.p_outlier;
    // wfpt.pyx:63-64: p_outlier outside [0,1] -> -inf for the whole set
    if (!((po >= 0) & (po <= 1))) {
        if (threadIdx.x == 0) {
            if (out_zero) { out_sum[bid] = 0.0; out_zero[bid] = 1; }
            else out_sum[bid] = -CUDART_INF;
            if (out_zero && done) final_out[p] = -CUDART_INF;
        }
        return;
    }
    Shared& S = shared();
    exptab_fetch(S.exptab, g_exp_table);
    int mode, kind, nz, nt, kl_max, sv_nonzero;
    SetConst sc;                            // per-thread copy: the loop reads these from registers
    {
        fast::Set set;
        fast::prepare(set, prm, o, kt);     // CTA-collective: fills fast::Tables (barrier inside)
        mode = set.mode; kind = set.kind; nz = set.nz; nt = set.nt; kl_max = set.kl_max; sv_nonzero = set.sv_nonzero;
        sc.a2 = set.a2; sc.inv_a2 = set.inv_a2; sc.sv2 = set.sv2; sc.nv2 = -set.v2;
        sc.pref_s = fast::kInvSqrt2Pi * set.inv_a2; sc.pref_l = kPi * set.inv_a2;
        sc.omp = __dsub_rn(1.0, po); sc.wp = __dmul_rn(w_outlier, po);
        sc.tlo = set.tlo;
        sc.ct = 1.0;
        if (threadIdx.x == 0) { S.ndefer = 0; S.a2 = sc.a2; }

        if (threadIdx.x < 2 && mode != fast::kModeZero && mode != fast::kModeGeneric) {
            const Tables& Tz = fast::tables();
            const int sg = threadIdx.x;
            S.pcz[0][sg] = sc.pref_s * Tz.cz[sg];
            S.pcz[1][sg] = sc.pref_l * Tz.cz[sg];
            S.gk[sg][0] = -2.0 * (Tz.z[sg][0].u[1] + 1.0);
            S.gk[sg][1] = -2.0 * (1.0 - Tz.z[sg][set.nz].u[1]);
            S.gk[sg][2] = -2.0 * Tz.zstep[sg];
            S.gk[sg][3] = 0.0;
        }
        for (int i = threadIdx.x; i < kWarps * (fast::kMaxNodes + 1); i += kBlock) {
            PieceCache e{0, 0, 0, 0};       // empty interval: the first round looks its pieces up
            (&S.piece[0][0])[i] = e;
        }
        if (mode != fast::kModeZero && mode != fast::kModeGeneric) {
            const Tables& T0 = fast::tables();
            for (int i = threadIdx.x; i < kt.n; i += kBlock)
                S.xthr_hi[i] = __double2hiint(__longlong_as_double(kt.thr[i]) * sc.a2);
            for (int i = threadIdx.x; i <= fast::kMaxNodes; i += kBlock) {
                TRec t{0.0, 0.0};
                if (i <= nt) { t.tn = T0.tn[i]; t.wt = T0.wts_t[i] * T0.ct; }   // ct == 1 when the t axis is not integrated
                S.trec[i] = t;
            }
        }
    }
    exptab_wait();
    __syncthreads();
    const Tables& T = fast::tables();

    const int64_t beg = (int64_t)s * chunk;
    const int64_t end = (beg + chunk < n) ? beg + chunk : n;
    int64_t b = beg;
    const bool head = (reinterpret_cast<uintptr_t>(x + b) & 15) && b < end;
    if (head) b++;
    const int npair = (int)((end - b) >> 1);
    const bool tail = (end > b) && ((end - b) & 1);
    const double2* x2 = reinterpret_cast<const double2*>(x + b);
    const int rounds = (npair + kBlock - 1) / kBlock + 1;   // + 1: the head / tail elements

    const unsigned tb = (unsigned)__cvta_generic_to_shared(&fast::tables());
    const unsigned etab = (unsigned)__cvta_generic_to_shared(S.exptab);
    LogProd lp;
    const bool usable = lean_set_ok(mode, kind);
    const bool zero_set = (mode == fast::kModeZero);
    const bool sv = sv_nonzero != 0;
    const bool z3 = (kind == 1);

    if (usable) {
        double2 nxt = make_double2(0.0, 0.0);          // the pair of the NEXT round is loaded a round ahead
        if ((int)threadIdx.x < npair) nxt = __ldg(x2 + threadIdx.x);
#pragma unroll 1
        for (int r = 0; r < rounds; r++) {
            double v0 = 0.0, v1 = 0.0;
            int cnt = 0;
            const int i = r * kBlock + threadIdx.x;
            const double2 xx = nxt;
            if (i + kBlock < npair) nxt = __ldg(x2 + i + kBlock);
            if (i < npair) {
                v0 = xx.x;
                v1 = xx.y;
                cnt = 2;
            } else if (r == rounds - 1) {
                if (head && threadIdx.x == 0) { v0 = x[beg]; cnt = 1; }
                if (tail && threadIdx.x == kBlock - 1) { v0 = x[end - 1]; cnt = 1; }
            }
            const int wcnt = __reduce_max_sync(kFull, cnt);
#pragma unroll 1
            for (int h = 0; h < wcnt; h++) {
                const double xv = h ? v1 : v0;
                const double ax = fabs(xv);
                const bool have = h < cnt;
                const bool live = have && !zero_set && !((ax - sc.tlo) < 0);   // pdf.pxi:112, the trial-dependent part
                double dens = 0.0;
                const int sign = xv > 0;
                int neg = 0, bad = 0;
                // (the three-node routine's own vote covers "somebody is live": the plain case pays one vote, not two)
                const bool plain = z3 && sv && nt == 2 && warp_density3<3, true>(ax, sign, live, S, sc, tb, etab, &dens, neg, bad);
                if (plain || __any_sync(kFull, live)) {
                    if (plain) {
                    } else if (z3 && sv) {
                        dens = warp_density<3, true>(ax, sign, live, T, S, sc, nt, tb, etab, neg, bad);
                    } else if (z3) dens = warp_density<3, false>(ax, sign, live, T, S, sc, nt, tb, etab, neg, bad);
                    else dens = sv ? warp_density<1, true>(ax, sign, live, T, S, sc, nt, tb, etab, neg, bad)
                                   : warp_density<1, false>(ax, sign, live, T, S, sc, nt, tb, etab, neg, bad);
                    // a negative truncated series anywhere makes pdf_sv's log(p) NaN (pdf.pxi:102); pdf (sv == 0) lets it through
                    if (sv && neg < 0) dens = CUDART_NAN;
                    if (!live) dens = 0.0;
                    if (live && bad) {
                        // an exp argument outside the branch-free kernel's range: the general evaluator takes this trial later
                        const int slot = atomicAdd(&S.ndefer, 1);
                        if (slot < kDeferCap) S.defer[slot] = xv;
                        continue;
                    }
                }
                if (have) lp.add(mix_outlier(dens, sc.omp, sc.wp));
            }
        }
    }
    __syncthreads();
    const int nd = S.ndefer;
    if (!usable || nd > kDeferCap) {
        // not a layout this kernel stages, or too many deferred trials: the whole slice through the general evaluator
        lp = LogProd();
#pragma unroll 1
        for (int64_t i = beg + threadIdx.x; i < end; i += kBlock)
            lp.add_slow(mix_outlier(general_density(x[i], kind, nz, nt, kl_max, mode, sc.tlo), sc.omp, sc.wp));
    } else {
#pragma unroll 1
        for (int q = threadIdx.x; q < nd; q += kBlock)
            lp.add_slow(mix_outlier(general_density(S.defer[q], kind, nz, nt, kl_max, mode, sc.tlo), sc.omp, sc.wp));
    }
    const double acc = lp.value();
    int any_zero;
    const double tot = block_sum_or<kBlock>(acc, lp.zero, &any_zero);
    if (threadIdx.x == 0) {
        if (out_zero) {
            out_sum[bid] = tot;
            out_zero[bid] = any_zero;
            if (done) {
                __threadfence();
                if (atomicAdd(&done[p], 1) == slices - 1) {
                    __threadfence();
                    double sum = 0.0;
                    int z = 0;
                    for (int k = 0; k < slices; k++) {
                        sum += __ldcg(&out_sum[p * slices + k]);
                        z |= __ldcg(&out_zero[p * slices + k]);
                    }
                    final_out[p] = z ? -CUDART_INF : sum;
                    done[p] = 0;
                }
            }
        } else {
            out_sum[bid] = any_zero ? -CUDART_INF : tot;   // wfpt.pyx:71-72
        }
    }
}

// ---- adaptive launches (use_adaptive != 0: what HDDM runs by default, likelihoods.py:105-113) -----------------------------------
// The first level of the reference's recursion (5 x 5 quarter points) runs in the warp-organised loop above; a trial whose
// recursion goes deeper is parked in shared memory (fast::Tables::queue) and evaluated by the refinement routines of
// wfpt_fast.cuh -- by the whole CTA whenever the lot is nearly full, and after the loop -- which split intervals and add
// results in the reference's order.
__device__ __forceinline__ void write_result(double tot, int any_zero, int64_t bid, int64_t p, int slices,
                                             double* __restrict__ out_sum, int* __restrict__ out_zero,
                                             int* __restrict__ done, double* __restrict__ final_out)
{
    if (out_zero) {
        out_sum[bid] = tot;
        out_zero[bid] = any_zero;
        if (done) {
            // whoever finishes last adds the set's partials up in slice order and leaves the counter at zero
            __threadfence();
            if (atomicAdd(&done[p], 1) == slices - 1) {
                __threadfence();
                double sum = 0.0;
                int z = 0;
                for (int k = 0; k < slices; k++) {
                    sum += __ldcg(&out_sum[p * slices + k]);
                    z |= __ldcg(&out_zero[p * slices + k]);
                }
                final_out[p] = z ? -CUDART_INF : sum;
                done[p] = 0;
            }
        }
    } else {
        out_sum[bid] = any_zero ? -CUDART_INF : tot;   // wfpt.pyx:71-72
    }
}

constexpr int kParkRounds = 3;   // rounds between two looks at the parking lot
static_assert(fast::kQueue > kParkRounds * 2 * kBlock, "parking lot smaller than what can arrive between two checks");

__global__ void __launch_bounds__(kBlock, 5) like_lean_adapt_kernel(const double* __restrict__ x, int64_t n,
                                                                   const double* __restrict__ params, int64_t P, Opts o,
                                                                   double w_outlier, int slices, int64_t chunk,
                                                                   double* __restrict__ part_sum, int* __restrict__ part_zero,
                                                                   double* __restrict__ park_x, int* __restrict__ park_n,
                                                                   int park_cap, const __grid_constant__ KTable kt)
{
    // Writes, per (set, slice) item: the sum of the log terms of the trials settled at the first level, the zero flag, and
    // the list of parked trials (park_x[item * park_cap ..], park_n[item]) for like_lean_refine_kernel, which adds the
    // parked trials' terms and produces the final sums.  Only a list that overflows its region is refined here.
    const int64_t bid = blockIdx.x;
    const int64_t p = bid / slices;
    const int s = (int)(bid - p * slices);
    if (p >= P) return;
    const double* row = params + 8 * p;
    Params prm{row[0], row[1], row[2], row[3], row[4], row[5], row[6], row[7]};
    const double po = prm.p_outlier;
    if (!((po >= 0) & (po <= 1))) {   // wfpt.pyx:63-64: the refine kernel turns the flag into -inf
        if (threadIdx.x == 0) { part_sum[bid] = 0.0; part_zero[bid] = 1; park_n[bid] = 0; }
        return;
    }
    Shared& S = shared();
    exptab_fetch(S.exptab, g_exp_table);
    int mode, kind, nt, kl_max, sv_nonzero;
    SetConst sc;
    {
        fast::Set set;
        fast::prepare(set, prm, o, kt);     // CTA-collective: fills fast::Tables (barrier inside), empties the parking lot
        mode = set.mode; kind = set.kind; nt = set.nt; kl_max = set.kl_max; sv_nonzero = set.sv_nonzero;
        sc.a2 = set.a2; sc.inv_a2 = set.inv_a2; sc.sv2 = set.sv2; sc.nv2 = -set.v2;
        sc.pref_s = fast::kInvSqrt2Pi * set.inv_a2; sc.pref_l = kPi * set.inv_a2;
        sc.omp = __dsub_rn(1.0, po); sc.wp = __dmul_rn(w_outlier, po);
        sc.tlo = set.tlo;
        sc.ct = 1.0;
        if (threadIdx.x == 0) S.a2 = sc.a2;

        if (threadIdx.x < 2 && mode != fast::kModeZero && mode != fast::kModeGeneric) {
            const Tables& Tz = fast::tables();
            const int sg = threadIdx.x;
            S.pcz[0][sg] = sc.pref_s * Tz.cz[sg];
            S.pcz[1][sg] = sc.pref_l * Tz.cz[sg];
            S.gk[sg][0] = -2.0 * (Tz.z[sg][0].u[1] + 1.0);
            S.gk[sg][1] = -2.0 * (1.0 - Tz.z[sg][set.nz].u[1]);
            S.gk[sg][2] = -2.0 * Tz.zstep[sg];
            S.gk[sg][3] = 0.0;
        }
        for (int i = threadIdx.x; i < kWarps * (fast::kMaxNodes + 1); i += kBlock) {
            PieceCache e{0, 0, 0, 0};
            (&S.piece[0][0])[i] = e;
        }
        if (mode == fast::kModeAdaptive || mode == fast::kModePoint) {
            const Tables& T0 = fast::tables();
            for (int i = threadIdx.x; i < kt.n; i += kBlock)
                S.xthr_hi[i] = __double2hiint(__longlong_as_double(kt.thr[i]) * sc.a2);
            for (int i = threadIdx.x; i <= fast::kMaxNodes; i += kBlock) {
                TRec t{0.0, 0.0};
                if (i <= nt) t.tn = T0.tn[i];
                S.trec[i] = t;
            }
        }
    }
    exptab_wait();
    __syncthreads();
    const Tables& T = fast::tables();
    const double simps_err = o.simps_err;
    const int depth_z = o.n_sz, depth_t = o.n_st;

    const int64_t beg = (int64_t)s * chunk;
    const int64_t end = (beg + chunk < n) ? beg + chunk : n;
    int64_t b = beg;
    const bool head = (reinterpret_cast<uintptr_t>(x + b) & 15) && b < end;
    if (head) b++;
    const int npair = (int)((end - b) >> 1);
    const bool tail = (end > b) && ((end - b) & 1);
    const double2* x2 = reinterpret_cast<const double2*>(x + b);
    const int rounds = (npair + kBlock - 1) / kBlock + 1;   // + 1: the head / tail elements

    const unsigned tb = (unsigned)__cvta_generic_to_shared(&fast::tables());
    const unsigned etab = (unsigned)__cvta_generic_to_shared(S.exptab);
    LogProd lp;
    const bool zero_set = (mode == fast::kModeZero);
    const bool usable = zero_set || mode == fast::kModePoint || mode == fast::kModeAdaptive;
    const bool sv = sv_nonzero != 0;
    const bool z5 = (kind == 3 || kind == 4);
    const bool parks = (mode == fast::kModeAdaptive || mode == fast::kModePoint);   // (a point set parks on exp range only)
    const bool point = (mode == fast::kModePoint);

    // parked trials: a handful -> one warp each (lanes share the trial's quadrature nodes); many -> one thread each
    auto drain = [&](int first, int parked) {   // lot entries first .. parked - 1
        if (point) {   // nothing to refine: the operation-for-operation evaluator takes the trial (library exp, every term)
#pragma unroll 1
            for (int q = first + threadIdx.x; q < parked; q += kBlock)
                lp.add_slow(mix_outlier(fast::generic_full_pdf(fast::parking_lot()[q], fast::tables().prm, fast::tables().o), sc.omp, sc.wp));
            return;
        }
#pragma unroll 1
        for (int q = first + threadIdx.x; q < parked; q += kBlock)
            lp.add_slow(mix_outlier(fast::adaptive_refine(fast::parking_lot()[q], &fast::tables(), kind, kl_max), sc.omp, sc.wp));
    };

    // the shared-memory lot goes to this item's region of the global list; what does not fit is refined here
    int flushed = 0;
    auto flush = [&](int parked) {
        const int room = park_cap - flushed;
        const int take = parked < room ? parked : room;
        double* dst = park_x + (size_t)bid * park_cap + flushed;
        for (int q = threadIdx.x; q < take; q += kBlock) dst[q] = fast::parking_lot()[q];
        flushed += take;
        if (take < parked) drain(take, parked);
    };
    if (usable) {
        double2 nxt = make_double2(0.0, 0.0);          // the pair of the NEXT round is loaded a round ahead
        if ((int)threadIdx.x < npair) nxt = __ldg(x2 + threadIdx.x);
#pragma unroll 1
        for (int r = 0; r < rounds; r++) {
            double v0 = 0.0, v1 = 0.0;
            int cnt = 0;
            const int i = r * kBlock + threadIdx.x;
            const double2 xx = nxt;
            if (i + kBlock < npair) nxt = __ldg(x2 + i + kBlock);
            if (i < npair) {
                v0 = xx.x;
                v1 = xx.y;
                cnt = 2;
            } else if (r == rounds - 1) {
                if (head && threadIdx.x == 0) { v0 = x[beg]; cnt = 1; }
                if (tail && threadIdx.x == kBlock - 1) { v0 = x[end - 1]; cnt = 1; }
            }
            const int wcnt = __reduce_max_sync(kFull, cnt);
#pragma unroll 1
            for (int h = 0; h < wcnt; h++) {
                const double xv = h ? v1 : v0;
                const double ax = fabs(xv);
                const bool have = h < cnt;
                const bool live = have && !zero_set && !((ax - sc.tlo) < 0);   // pdf.pxi:112, the trial-dependent part
                double dens = 0.0;
                if (__any_sync(kFull, live)) {
                    const int sign = xv > 0;
                    int neg = 0, bad = 0;
                    bool deeper = false;
                    if (z5) dens = sv ? warp_density_adapt<5, true>(ax, sign, live, T, S, sc, nt, tb, etab, simps_err, depth_z, depth_t, neg, bad, deeper)
                                      : warp_density_adapt<5, false>(ax, sign, live, T, S, sc, nt, tb, etab, simps_err, depth_z, depth_t, neg, bad, deeper);
                    else dens = sv ? warp_density_adapt<1, true>(ax, sign, live, T, S, sc, nt, tb, etab, simps_err, depth_z, depth_t, neg, bad, deeper)
                                   : warp_density_adapt<1, false>(ax, sign, live, T, S, sc, nt, tb, etab, simps_err, depth_z, depth_t, neg, bad, deeper);
                    // a negative truncated series anywhere makes pdf_sv's log(p) NaN (pdf.pxi:102); pdf (sv == 0) lets it through
                    if (sv && neg < 0) dens = CUDART_NAN;
                    if (!live) dens = 0.0;
                    if (live && (deeper || bad)) {
                        // below the first level, or an exp argument outside the branch-free kernel's range: the refinement
                        // routines take the trial (the lot is drained before it can overflow, see below)
                        const int slot = atomicAdd(&fast::tables().qcount, 1);
                        if (slot >= fast::kQueue) __trap();
                        fast::parking_lot()[slot] = xv;
                        continue;
                    }
                }
                if (have) lp.add(mix_outlier(dens, sc.omp, sc.wp));
            }
            if (parks && (r % kParkRounds) == kParkRounds - 1) {   // uniform over the CTA
                __syncthreads();
                const int parked = fast::tables().qcount;
                if (parked > fast::kQueue - kParkRounds * 2 * kBlock) {
                    flush(parked);
                    __syncthreads();
                    if (threadIdx.x == 0) fast::tables().qcount = 0;
                    __syncthreads();
                }
            }
        }
        if (parks) {
            __syncthreads();
            flush(fast::tables().qcount);
        }
    } else {
        // a layout the tables do not cover: every trial through the general evaluator
#pragma unroll 1
        for (int64_t i = beg + threadIdx.x; i < end; i += kBlock)
            lp.add_slow(mix_outlier(fast::generic_full_pdf(x[i], fast::tables().prm, fast::tables().o), sc.omp, sc.wp));
    }
    const double acc = lp.value();
    int any_zero;
    const double tot = block_sum_or<kBlock>(acc, lp.zero, &any_zero);
    if (threadIdx.x == 0) {
        part_sum[bid] = tot;
        part_zero[bid] = any_zero;
        park_n[bid] = flushed;
    }
}

// Second kernel of an adaptive launch: one CTA per (set, slice) item refines the trials the first kernel parked for it
// (integrate.pxi:72-206 below the first level; the routines of wfpt_fast.cuh: one warp per trial for a handful, one thread
// per trial otherwise), adds their log terms to the item's first-level sum in a fixed order and writes the result.  Keeping
// the refinement out of the first kernel keeps both instruction working sets small: run inside it, the two code paths
// evicted each other from the instruction cache (7 of 13 stalled warps per issue cycle waiting for instructions on the
// whole-box grid).
__global__ void __launch_bounds__(kBlock, 4) like_lean_refine_kernel(const double* __restrict__ params, int64_t P, Opts o,
                                                                    double w_outlier, int slices,
                                                                    double* __restrict__ part_sum, int* __restrict__ part_zero,
                                                                    const double* __restrict__ park_x,
                                                                    const int* __restrict__ park_n, int park_cap,
                                                                    int* __restrict__ done, double* __restrict__ final_out,
                                                                    const __grid_constant__ KTable kt)
{
    const int64_t bid = blockIdx.x;
    const int64_t p = bid / slices;
    if (p >= P) return;
    const int cnt = park_n[bid];
    double tot = part_sum[bid];
    int any_zero = part_zero[bid];
    if (cnt > 0) {      // uniform over the CTA
        const double* row = params + 8 * p;
        Params prm{row[0], row[1], row[2], row[3], row[4], row[5], row[6], row[7]};
        const double po = prm.p_outlier;
        const double omp = __dsub_rn(1.0, po), wp = __dmul_rn(w_outlier, po);
        fast::Set set;
        fast::prepare(set, prm, o, kt);
        const int kind = set.kind, kl_max = set.kl_max;
        const double* list = park_x + (size_t)bid * park_cap;
        LogProd lp;
        if (set.mode == fast::kModePoint) {
#pragma unroll 1
            for (int q = threadIdx.x; q < cnt; q += kBlock)
                lp.add_slow(mix_outlier(fast::generic_full_pdf(list[q], fast::tables().prm, fast::tables().o), omp, wp));
        } else if (cnt <= 32 && kind != 4 && o.n_st == 2) {
#pragma unroll 1
            for (int q = threadIdx.x >> 5; q < cnt; q += kWarps) {
                const double p0 = mix_outlier(fast::refine_coop(list[q], &fast::tables(), kind, kl_max), omp, wp);
                if ((threadIdx.x & 31) == 0) lp.add_slow(p0);
            }
        } else {
#pragma unroll 1
            for (int q = threadIdx.x; q < cnt; q += kBlock)
                lp.add_slow(mix_outlier(fast::adaptive_refine(list[q], &fast::tables(), kind, kl_max), omp, wp));
        }
        int z2;
        const double add = block_sum_or<kBlock>(lp.value(), lp.zero, &z2);
        tot += add;
        any_zero |= z2;
    }
    if (threadIdx.x == 0) {
        if (slices > 1) write_result(tot, any_zero, bid, p, slices, part_sum, part_zero, done, final_out);
        else final_out[p] = any_zero ? -CUDART_INF : tot;
    }
}

}  // namespace lean

// Which fixed-grid launches the lean kernel takes (decided from the options alone; per-set layouts inside are then always
// ones it stages: no integration, t only, or n_sz = 2 with any n_st).
bool lean_applies(const BatchArgs& a, const KTable& kt)
{
    return !a.o.use_adaptive && !a.seg && !a.one_each && kt.n >= 0 && kt.kl_max <= 4 && kt.ks_max <= 4 && a.o.n_sz == 2 &&
           a.o.n_st >= 1 && a.o.n_st < fast::kMaxNodes;
}

// 2^(j/2048) rounded correctly (long double exp2l), uploaded once per device.
static cudaError_t ensure_exp_table()
{
    static std::mutex mu;
    static bool ready[64] = {false};
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return e;
    if (dev < 0 || dev >= 64) return cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lk(mu);
    if (ready[dev]) return cudaSuccess;
    static double host[lean::kExpN];
    for (int j = 0; j < lean::kExpN; j++) {
        const double v = (double)exp2l((long double)j / (long double)lean::kExpN);   // correctly rounded 2^(j/N)
        uint64_t bits;
        memcpy(&bits, &v, 8);
        bits -= (uint64_t)j << (32 + 9);                                             // see expv
        memcpy(&host[j], &bits, 8);
    }
    e = cudaMemcpyToSymbol(lean::g_exp_table, host, sizeof(host));
    if (e == cudaSuccess) ready[dev] = true;
    return e;
}

bool lean_adapt_applies(const BatchArgs& a, const KTable& kt)
{
    return a.o.use_adaptive && !a.seg && !a.one_each && a.park_x && a.park_n && a.part_sum && a.part_zero && kt.n >= 0 &&
           kt.kl_max <= 4 && kt.ks_max <= 4;
}

cudaError_t launch_like_lean(const BatchArgs& a, const KTable& kt)
{
    {
        const cudaError_t e = ensure_exp_table();
        if (e != cudaSuccess) return e;
    }
    const int64_t nblk = a.P * (int64_t)a.slices;
    if (nblk == 0) return cudaSuccess;
    if (a.o.use_adaptive) {
        lean::like_lean_adapt_kernel<<<(unsigned)nblk, kBlock, 0, a.stream>>>(
            a.x, a.n, a.params, a.P, a.o, a.w_outlier, a.slices, a.chunk, a.part_sum, a.part_zero, a.park_x, a.park_n,
            a.park_cap, kt);
        cudaError_t e = cudaGetLastError();
        if (e != cudaSuccess) return e;
        lean::like_lean_refine_kernel<<<(unsigned)nblk, kBlock, 0, a.stream>>>(
            a.params, a.P, a.o, a.w_outlier, a.slices, a.part_sum, a.part_zero, a.park_x, a.park_n, a.park_cap,
            a.slices > 1 ? a.done : nullptr, a.out, kt);
        count_launch(1);
        return cudaGetLastError();
    }
    lean::like_lean_kernel<<<(unsigned)nblk, kBlock, 0, a.stream>>>(
        a.x, a.n, a.params, a.P, a.o, a.w_outlier, a.slices, a.chunk, a.slices > 1 ? a.part_sum : a.out,
        a.slices > 1 ? a.part_zero : nullptr, a.slices > 1 ? a.done : nullptr, a.out, kt);
    return cudaGetLastError();
}

}  // namespace wfpt
