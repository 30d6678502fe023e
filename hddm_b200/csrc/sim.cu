// Euler-Maruyama sequential-sampling simulators (one particle = one first-passage walk) for sm_100a.
//
// What is simulated (paths relative to /root/reference/src/cddm_data_simulation.pyx; dispatch and theta column
// order from /root/reference/graveyard/basic_simulator_old.py:168-480 and hddm/model_config.py):
//   WFPT_SIM_DDM              ddm                 :294-375   bounds [0, a],  start z*a, choice 0/1
//   WFPT_SIM_FULL_DDM_BASE    full_ddm_hddm_base  :174-284   bounds [0, a],  sv/sz/st draws in absolute units
//   WFPT_SIM_DDM_FLEXBOUND    ddm_flexbound       :454-553   symmetric +-a (constant bound), choice sign(y)
//   WFPT_SIM_ANGLE            ddm_flexbound + angle bound  b(t) = a + t * (-sin(th)/cos(th))   (additive)
//   WFPT_SIM_WEIBULL          ddm_flexbound + Weibull bound b(t) = a * exp(-(t/beta)^alpha)    (multiplicative)
//   WFPT_SIM_FULL_DDM         full_ddm            :747-869
//   WFPT_SIM_DDM_SDV          ddm_sdv             :876-994
//   WFPT_SIM_ORNSTEIN         ornstein_uhlenbeck  :1119-1219
//   WFPT_SIM_LEVY             levy_flexbound      :642-741   alpha-stable increments (Chambers-Mallows-Stuck, :28-40)
//   WFPT_SIM_GAMMA_DRIFT(_ANGLE), WFPT_SIM_DS_CONFLICT(_ANGLE)   ddm_flex :1001-1113 with the time-varying drifts of
//                             graveyard/drift_functions.py:11-103 added to v, evaluated in registers at every step
//   WFPT_SIM_DDM_COV          ddm_cov             :384-449   ddm with choice = -sign(y)
//
// Arithmetic follows the reference: particle state, time and bounds are float32; time is accumulated by repeated
// float32 addition of delta_t (so RTs land on the reference's grid of values); the increment is
// y += (v*dt) + (sqrt(dt)*s*N) with each product and sum rounded on its own.  The reference's RNG (C rand() + polar
// method, cyclic buffer) is replaced by counter-based Philox4x32 (seven rounds, see below): key = seed, counter = (particle, set, block), so
// particle p of parameter set k sees the same stream whatever the launch geometry or the number of GPUs.  Parity is
// distributional ("parity unpinned" for simulator numerics).
//
// Mapping: one CTA works on one parameter set at a time (its constants in registers, its RT/choice histogram in
// shared memory, built with shared-memory atomics and flushed once per item); each thread runs one particle after
// another without waiting for its warp neighbours (a finished lane picks up its next particle at the next Philox
// block boundary, at most three sub-steps later); bounds are evaluated in registers.
#include <cstdio>
#include <string>

#include "../../include/wfpt_b200.h"
#include "common.cuh"

namespace wfpt {
namespace sim {

constexpr int kSimBlock = 256;
constexpr int kMaxBins = 4096;

struct Args {
    const float* theta;  // [n_sets][n_par]
    int n_par;
    int n_sets;
    long long n_samples;       // particles per set in this launch
    long long sample_offset;   // global index of this launch's first particle (multi-GPU particle sharding)
    int set_offset = 0, set_stride = 1;   // global index of local set k is set_offset + k * set_stride (multi-GPU set sharding):
                                          // the Philox counter carries the GLOBAL set index, so a shard reproduces the unsharded run
    long long out_stride_samples;  // rows of the (n_samples_total, n_sets) output arrays written by this launch start at 0
    float dt, max_t, s;
    unsigned long long seed;
    float* rts;                // [n_samples][n_sets] or nullptr
    int* choices;              // same
    unsigned int* hist;        // [n_sets][nbins][2] or nullptr
    int nbins;
    double hist_step;          // max_t_hist / (nbins - 1)
    float hist_inv_step;       // its reciprocal in float32: first guess of the bin, corrected against the exact edges
    double hist_max_t;
    unsigned long long* steps; // optional: total Euler steps
    float* traj;               // optional: trajectory of (set 0, particle 0): [int(max_t/dt)+1]
    int traj_len;
    long long chunk;           // particles per work item
    long long items_per_set;
    unsigned int key_a[10], key_b[10];   // Philox round keys of `seed`
    void set_seed(unsigned long long sd)
    {
        seed = sd;
        for (int r = 0; r < 10; r++) {
            key_a[r] = (unsigned int)sd + (unsigned int)r * 0x9E3779B9u;
            key_b[r] = (unsigned int)(sd >> 32) + (unsigned int)r * 0xBB67AE85u;
        }
    }
};

// ---- Philox4x32 ----------------------------------------------------------------------------------------------------------
// The ten round keys (the Weyl sequence k + r*W) depend on the seed only.  The host computes them into the kernel
// arguments (Args::set_seed), i.e. into the constant bank, where the round's XOR reads them as an operand: no
// instruction and no register is spent on them.
struct RoundKeys {
    const unsigned int *a, *b;
    __device__ __forceinline__ RoundKeys(const Args& A) : a(A.key_a), b(A.key_b) {}
};

__device__ __forceinline__ uint4 philox4x32(unsigned int c0, unsigned int c1, unsigned int c2, unsigned int c3,
                                            const RoundKeys& k)
{
#ifndef WFPT_PHILOX_ROUNDS
#define WFPT_PHILOX_ROUNDS 7    // Philox4x32-7: Random123's Crush-resistant round count (-10 is its conservative default).
                                // Measured against -10 on the config-5 launch: +13.7 % Euler steps/s, choice probability, RT
                                // moments, lag-1 correlation and the KS / CI suite of tests/test_gpu_sim.py unchanged
#endif
#pragma unroll
    for (int r = 0; r < WFPT_PHILOX_ROUNDS; r++) {
        // (separate high / low multiplies: the 64-bit product form costs two extra instructions per multiply here)
        const unsigned int h0 = __umulhi(0xD2511F53u, c0), l0 = 0xD2511F53u * c0;
        const unsigned int h1 = __umulhi(0xCD9E8D57u, c2), l1 = 0xCD9E8D57u * c2;
        c0 = h1 ^ c1 ^ k.a[r];
        c2 = h0 ^ c3 ^ k.b[r];
        c1 = l1;
        c3 = l0;
    }
    return make_uint4(c0, c1, c2, c3);
}

__device__ __forceinline__ float u01(unsigned int x) { return ((float)x + 0.5f) * 2.3283064365386963e-10f; }  // (0, 1)

// Box-Muller radius sqrt(-2 ln u) for u = (x + 0.5) 2^-32 in (0, 1): one lg2 and one rsqrt on the SFU, no special cases
// (u is never 0, 1 or subnormal, so the guarded library forms of log and sqrt are not needed).
__device__ __forceinline__ float bm_radius(unsigned int x)
{
    const float u = fmaf((float)x, 2.3283064365386963e-10f, 1.1641532182693481e-10f);
    float l2;
    asm("lg2.approx.ftz.f32 %0, %1;" : "=f"(l2) : "f"(u));
    const float q = fmaxf(l2 * -1.3862943611198906f, 1e-30f);   // -2 ln u  (> 0; the floor only guards u rounding to 1)
    float rs;
    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rs) : "f"(q));
    return q * rs;
}

// Four standard normals from one Philox block (two Box-Muller pairs on the special-function unit).
__device__ __forceinline__ float4 normals4(const uint4 r)
{
    const float r0 = bm_radius(r.x), r1 = bm_radius(r.z);
    const float a0 = fmaf((float)r.y, 1.4629180792671596e-09f, 7.314590396335798e-10f);   // 2 pi (y + 0.5) 2^-32
    const float a1 = fmaf((float)r.w, 1.4629180792671596e-09f, 7.314590396335798e-10f);
    return make_float4(r0 * __cosf(a0), r0 * __sinf(a0), r1 * __cosf(a1), r1 * __sinf(a1));
}

// Chambers-Mallows-Stuck draw as the reference writes it (cddm_data_simulation.pyx:28-40)
__device__ __forceinline__ float stable_draw(float alpha, float U, float U2)
{
    const float u = 3.14159265358979f * (U - 0.5f);
    const float w = -__logf(U2);
    if (alpha == 1.0f) return __tanf(u);
    return (__sinf(alpha * u) / __powf(__cosf(u), 1.0f / alpha)) * __powf(__cosf(u - alpha * u) / w, (1.0f - alpha) / alpha);
}

// ---- per-model constants --------------------------------------------------------------------------------------------------
struct SetConst {
    float v, a, z, t;
    float p4, p5, p6;      // model-specific extras (theta columns 4..6)
    float vdt;             // v * dt
    float slope;           // angle: -sin(th)/cos(th)
    float sdt;             // angle: slope * dt (bound decrement per step)
    float inv_beta, alpha; // weibull
    float g;               // ornstein
    float dt_alpha;        // levy: s * dt^(1/alpha)
    // time-varying drift (ddm_flex): gamma: c/div, shape-1, -1/scale ; conflict: (init_t - fix_t)*coh_t, fix_t*coh_t, init_d*coh_d, slopes
    float d0, d1, d2, d3, d4;
};

template <int MODEL>
struct Traits {
    static constexpr bool zero_a = (MODEL == WFPT_SIM_DDM || MODEL == WFPT_SIM_FULL_DDM_BASE || MODEL == WFPT_SIM_DDM_COV);  // bounds [0, a]
    static constexpr bool angle = (MODEL == WFPT_SIM_ANGLE || MODEL == WFPT_SIM_GAMMA_DRIFT_ANGLE || MODEL == WFPT_SIM_DS_CONFLICT_ANGLE);
    static constexpr bool moving = angle || (MODEL == WFPT_SIM_WEIBULL);
    static constexpr bool full = (MODEL == WFPT_SIM_FULL_DDM_BASE || MODEL == WFPT_SIM_FULL_DDM);
    static constexpr bool sdv = (MODEL == WFPT_SIM_DDM_SDV);
    static constexpr bool gamma = (MODEL == WFPT_SIM_GAMMA_DRIFT || MODEL == WFPT_SIM_GAMMA_DRIFT_ANGLE);
    static constexpr bool conflict = (MODEL == WFPT_SIM_DS_CONFLICT || MODEL == WFPT_SIM_DS_CONFLICT_ANGLE);
    static constexpr bool tdrift = gamma || conflict;   // drift is a function of time
};

template <int MODEL>
__device__ __forceinline__ void load_set(SetConst& c, const float* __restrict__ th, const Args& A)
{
    c.v = th[0];
    c.a = th[1];
    c.z = th[2];
    c.t = th[3];
    c.p4 = A.n_par > 4 ? th[4] : 0.f;
    c.p5 = A.n_par > 5 ? th[5] : 0.f;
    c.p6 = A.n_par > 6 ? th[6] : 0.f;
    c.slope = 0.f; c.sdt = 0.f; c.inv_beta = 0.f; c.alpha = 0.f; c.g = 0.f; c.dt_alpha = 0.f;
    c.d0 = c.d1 = c.d2 = c.d3 = c.d4 = 0.f;
    float ang = c.p4;                                                               // angle model: v, a, z, t, theta
    if (MODEL == WFPT_SIM_WEIBULL) { c.alpha = c.p4; c.inv_beta = 1.0f / c.p5; }
    if (MODEL == WFPT_SIM_ORNSTEIN) { c.g = th[3]; c.t = th[4]; }             // theta = v, a, z, g, t
    if (MODEL == WFPT_SIM_LEVY) { c.alpha = th[3]; c.t = th[4]; c.dt_alpha = A.s * powf(A.dt, 1.0f / th[3]); }  // v, a, z, alpha, t
    if (Traits<MODEL>::gamma) {
        // theta = v, a, z, t, [theta,] shape, scale, c   (basic_simulator_old.py:343-377)
        const int o = (MODEL == WFPT_SIM_GAMMA_DRIFT_ANGLE) ? 5 : 4;
        const float shape = th[o], scale = th[o + 1], cc = th[o + 2];
        // c * t^(shape-1) exp(-t/scale) / ((shape-1)^(shape-1) scale^(shape-1) exp(-(shape-1)))   (drift_functions.py:30-36)
        const float div = powf(shape - 1.f, shape - 1.f) * powf(scale, shape - 1.f) * expf(-(shape - 1.f));
        c.d0 = cc / div;
        c.d1 = shape - 1.f;
        c.d2 = -1.0f / scale;
    }
    if (Traits<MODEL>::conflict) {
        // theta = a, z, t, tinit, dinit, tslope, dslope, tfixedp, tcoh, dcoh [, angle]; v = 0   (:379-425)
        c.v = 0.f; c.a = th[0]; c.z = th[1]; c.t = th[2];
        const float tinit = th[3], dinit = th[4], tslope = th[5], dslope = th[6], tfix = th[7], tcoh = th[8], dcoh = th[9];
        // v(t) = ((tinit - tfix) e^{-tslope t} + tfix) tcoh + (dinit e^{-dslope t}) dcoh   (drift_functions.py:39-103)
        c.d0 = (tinit - tfix) * tcoh;
        c.d1 = -tslope;
        c.d2 = tfix * tcoh;
        c.d3 = dinit * dcoh;
        c.d4 = -dslope;
        ang = A.n_par > 10 ? th[10] : 0.f;
    }
    if (Traits<MODEL>::angle) { c.slope = -sinf(ang) / cosf(ang); c.sdt = c.slope * A.dt; }
    c.vdt = __fmul_rn(c.v, A.dt);
}

// Time-varying part of the drift at time ts (ddm_flex: drift[ix] = v + f(t_s[ix]))
template <int MODEL>
__device__ __forceinline__ float drift_at(const SetConst& c, float ts)
{
    if (Traits<MODEL>::gamma) {
        const float pw = (ts > 0.f) ? exp2f(c.d1 * __log2f(ts)) : (c.d1 == 0.f ? 1.f : 0.f);
        return c.v + c.d0 * pw * __expf(c.d2 * ts);
    }
    if (Traits<MODEL>::conflict) return c.d0 * __expf(c.d1 * ts) + c.d2 + c.d3 * __expf(c.d4 * ts);
    return c.v;
}

// Upper bound after `fix` steps (time ts = fix*dt); the lower bound is -b for the flexbound family and 0 for the [0,a]
// family.  The reference tabulates b on t_s = float32(ix*dt); here it is evaluated in registers, the angle bound as one
// FMA a + fix*(slope*dt) (differs from the table by at most an ulp of b -- far below the sqrt(dt) step noise).
template <int MODEL>
__device__ __forceinline__ float bound_at(const SetConst& c, float fix, float dt)
{
    if (Traits<MODEL>::angle) return fmaf(fix, c.sdt, c.a);
    if (MODEL == WFPT_SIM_WEIBULL) {
        // a * exp(-(ts/beta)^alpha);  ts == 0 -> a
        const float r = (fix * dt) * c.inv_beta;
        const float pw = (r > 0.f) ? exp2f(c.alpha * __log2f(r)) : 0.f;
        return c.a * __expf(-pw);
    }
    return c.a;
}

// ---- the kernel ------------------------------------------------------------------------------------------------------------
// Stream layout of particle (set, p): Philox key = seed, counter = (p_lo, p_hi, set, block).  Block 0xFFFFFFFF feeds the
// start-of-particle draws (drift, non-decision time and starting point variability); block j = 0, 1, ... feeds Euler
// steps 4j .. 4j+3 (one normal each; Levy: steps 2j, 2j+1, two uniforms each).  All lanes of a warp fetch their blocks
// at the same loop iteration: a lane whose particle ends mid-block idles for at most three sub-steps and starts its
// next particle at the next block boundary, so the Philox rounds and the Box-Muller transform never run diverged.
// TRACE: a one-particle instantiation that also writes the path of (set 0, particle 0) -- kept out of the production loop.
template <int MODEL, bool TRACE>
__global__ void __launch_bounds__(kSimBlock) sim_kernel(const Args A)
{
    extern __shared__ unsigned int s_hist[];  // [nbins][2] when A.hist
    __shared__ int s_next;                    // next undistributed particle of the current item
    const int nh = A.hist ? A.nbins * 2 : 0;
    const float sqrt_st = __fmul_rn(sqrtf(A.dt), A.s);
    const RoundKeys rk(A);
    unsigned long long my_steps = 0;
    const long long n_items = (long long)A.n_sets * A.items_per_set;

    for (long long item = blockIdx.x; item < n_items; item += gridDim.x) {
        const int set = (int)(item / A.items_per_set);
        const long long beg = (item - (long long)set * A.items_per_set) * A.chunk;
        const long long end = (beg + A.chunk < A.n_samples) ? beg + A.chunk : A.n_samples;
        for (int i = threadIdx.x; i < nh; i += kSimBlock) s_hist[i] = 0u;
        __syncthreads();

        SetConst c;
        load_set<MODEL>(c, A.theta + (size_t)set * A.n_par, A);
        const unsigned int gset = (unsigned int)(A.set_offset + set * A.set_stride);

        // particles are handed out one at a time from a shared counter: a lane that finishes takes the next one, so the
        // lanes of the CTA run dry together (results do not depend on who simulates which particle)
        if (threadIdx.x == 0) s_next = kSimBlock;
        __syncthreads();
        long long p = beg + threadIdx.x;   // particle (local index within this launch)
        bool active = p < end;
        bool walking = false;
        float y = 0.f, tp = 0.f, fix = 0.f, drift = c.vdt, t_nd = c.t, b = c.a;
        unsigned int blk = 0, plo = 0, phi = 0;
        while (active) {
            if (!walking) {
                // ---- start a particle ----
                const unsigned long long gp = (unsigned long long)(p + A.sample_offset);
                plo = (unsigned int)gp;
                phi = (unsigned int)(gp >> 32);
                blk = 0;
                tp = 0.f;
                fix = 0.f;
                drift = c.vdt;
                t_nd = c.t;
                b = bound_at<MODEL>(c, 0.f, A.dt);
                if (Traits<MODEL>::zero_a) y = __fmul_rn(c.z, c.a);
                else y = __fadd_rn(-b, __fmul_rn(__fmul_rn(c.z, 2.f), b));
                if (Traits<MODEL>::full || Traits<MODEL>::sdv) {
                    const uint4 r = philox4x32(plo, phi, gset, 0xFFFFFFFFu, rk);
                    const float g = sqrtf(-2.0f * __logf(u01(r.x))) * __cosf(6.283185307179586f * u01(r.y));
                    if (Traits<MODEL>::full) {
                        // theta = v, a, z, t, sz, sv, st  (:228-235, :812-820)
                        drift = __fmul_rn(__fadd_rn(c.v, __fmul_rn(c.p5, g)), A.dt);
                        t_nd = __fadd_rn(c.t, __fmul_rn(__fmul_rn(2.f, u01(r.z) - 0.5f), c.p6));
                        y = __fadd_rn(y, __fmul_rn(__fmul_rn(2.f, u01(r.w) - 0.5f), c.p4));
                    } else {
                        drift = __fmul_rn(__fadd_rn(c.v, __fmul_rn(c.p4, g)), A.dt);  // theta = v, a, z, t, sv
                    }
                }
                if (TRACE) A.traj[0] = y;
                walking = true;
            }
            // ---- one Philox block for everybody, then up to four sub-steps ----
            const uint4 r = philox4x32(plo, phi, gset, blk++, rk);
            float4 nn;
            if (MODEL == WFPT_SIM_LEVY) {
                nn.x = stable_draw(c.alpha, u01(r.x), u01(r.y));
                nn.y = stable_draw(c.alpha, u01(r.z), u01(r.w));
                nn.z = nn.w = 0.f;
            } else {
                nn = normals4(r);
            }
            constexpr int kSub = (MODEL == WFPT_SIM_LEVY) ? 2 : 4;
            bool done = false;
#pragma unroll
            for (int j = 0; j < kSub; j++) {
                const float lo = Traits<MODEL>::zero_a ? 0.f : -b;
                const bool in = (!done) & (y >= lo) & (y <= b) & (tp <= A.max_t);
                done |= !in;
                // ---- one Euler step (predicated: lanes that have left the bounds keep their state) ----
                const float nj = (j == 0) ? nn.x : (j == 1) ? nn.y : (j == 2) ? nn.z : nn.w;
                const float inc = (MODEL == WFPT_SIM_LEVY) ? __fmul_rn(c.dt_alpha, nj) : __fmul_rn(sqrt_st, nj);
                float det = drift;
                if (MODEL == WFPT_SIM_ORNSTEIN) det = __fmul_rn(__fsub_rn(c.v, __fmul_rn(c.g, y)), A.dt);
                if (Traits<MODEL>::tdrift) det = __fmul_rn(drift_at<MODEL>(c, __fmul_rn(fix, A.dt)), A.dt);  // drift[ix] * dt
                // selects, not branches: the four sub-steps stay one basic block
                const float y_new = __fadd_rn(y, __fadd_rn(det, inc));
                const float tp_new = __fadd_rn(tp, A.dt);
                const float fix_new = fix + 1.0f;
                y = in ? y_new : y;
                tp = in ? tp_new : tp;
                fix = in ? fix_new : fix;
                if (Traits<MODEL>::moving) {
                    const float b_new = bound_at<MODEL>(c, fix_new, A.dt);
                    b = in ? b_new : b;
                }
                if (TRACE && in && (int)fix < A.traj_len) A.traj[(int)fix] = y;
            }
            if (done) {
                // ---- record and take the next particle ----
                my_steps += (unsigned int)fix;
                const float rt = __fadd_rn(tp, t_nd);
                int choice;
                if (MODEL == WFPT_SIM_DDM_COV) choice = (y < 0.f) - (y > 0.f);   // -sign(y)
                else if (Traits<MODEL>::zero_a) choice = (y < 0.f) ? 0 : 1;
                else choice = (y > 0.f) - (y < 0.f);
                if (A.rts) {
                    const size_t o = (size_t)p * A.n_sets + set;
                    A.rts[o] = rt;
                    A.choices[o] = choice;
                }
                if (A.hist) {
                    const int col = (Traits<MODEL>::zero_a && MODEL != WFPT_SIM_DDM_COV) ? choice : (choice > 0 ? 1 : (choice < 0 ? 0 : -1));
                    if (col >= 0) {
                        // np.histogram on edges linspace(0, max_t, nbins) + [inf] (basic_simulator_old.py:99-114)
                        const double rd = (double)rt;
                        int bin = (int)(rt * A.hist_inv_step);   // within one bin of the truth; settled below
                        if (bin > A.nbins - 1) bin = A.nbins - 1;
                        if (bin < 0) bin = 0;
                        const double e_lo = (bin == A.nbins - 1) ? A.hist_max_t : bin * A.hist_step;
                        if (rd < e_lo && bin > 0) bin--;
                        else if (bin < A.nbins - 1) {
                            const double e_hi = (bin + 1 == A.nbins - 1) ? A.hist_max_t : (bin + 1) * A.hist_step;
                            if (rd >= e_hi) bin++;
                        }
                        if (rd >= 0.0) atomicAdd(&s_hist[bin * 2 + col], 1u);
                    }
                }
                walking = false;
                p = beg + (long long)atomicAdd(&s_next, 1);
                active = p < end;
            }
        }
        __syncthreads();
        if (A.hist) {
            unsigned int* g = A.hist + (size_t)set * nh;
            for (int i = threadIdx.x; i < nh; i += kSimBlock) {
                const unsigned int v = s_hist[i];
                if (v) atomicAdd(&g[i], v);
            }
        }
        __syncthreads();
    }
    if (A.steps) {
        // CTA sum of the step counts, one global atomic per CTA
        __shared__ unsigned long long s_steps[kSimBlock / 32];
        for (int off = 16; off > 0; off >>= 1) my_steps += __shfl_down_sync(0xffffffffu, my_steps, off);
        if ((threadIdx.x & 31) == 0) s_steps[threadIdx.x >> 5] = my_steps;
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned long long tot = 0;
            for (int i = 0; i < kSimBlock / 32; i++) tot += s_steps[i];
            atomicAdd(A.steps, tot);
        }
    }
}

// ---- multi-accumulator models ---------------------------------------------------------------------------------------------
// race_model (cddm_data_simulation.pyx:1253-1374) and lca (:1380-1518): NP accumulators start at z_j * b(0) and run
// until one exceeds the (common, possibly collapsing) bound b(t) or time runs out; choice = index of the largest.
//   race: p_j += v_j dt + s sqrt(dt) N
//   lca : p_j += (v_j - g p_j - b_inh (sum_old - p_j)) dt + s sqrt(dt) N, then p_j = max(0, p_j)
// theta (canonical layout built by the host shim): v[NP], a, z[NP], g, b_inh, t, angle  -> 2 NP + 5 floats per set.
// One Philox block per Euler step (NP <= 4 normals), so the lanes of a warp are always in phase.
template <int NP, bool LCA, bool ANGLE>
__global__ void __launch_bounds__(kSimBlock) race_kernel(const Args A)
{
    extern __shared__ unsigned int s_hist[];  // [nbins][NP] when A.hist
    __shared__ int s_next;
    const int nh = A.hist ? A.nbins * NP : 0;
    const float sqrt_st = __fmul_rn(sqrtf(A.dt), A.s);
    const RoundKeys rk(A);
    unsigned long long my_steps = 0;
    const long long item = blockIdx.x;
    const int set = (int)(item / A.items_per_set);
    const long long beg = (item - (long long)set * A.items_per_set) * A.chunk;
    const long long end = (beg + A.chunk < A.n_samples) ? beg + A.chunk : A.n_samples;
    for (int i = threadIdx.x; i < nh; i += kSimBlock) s_hist[i] = 0u;
    if (threadIdx.x == 0) s_next = kSimBlock;
    __syncthreads();

    const float* th = A.theta + (size_t)set * A.n_par;
    float v[NP], z[NP], vdt[NP];
#pragma unroll
    for (int j = 0; j < NP; j++) {
        v[j] = th[j];
        z[j] = th[NP + 1 + j];
        vdt[j] = __fmul_rn(v[j], A.dt);
    }
    const float a = th[NP], g = th[2 * NP + 1], binh = th[2 * NP + 2], t_nd = th[2 * NP + 3];
    const float sdt = ANGLE ? (-sinf(th[2 * NP + 4]) / cosf(th[2 * NP + 4])) * A.dt : 0.f;

    long long p = beg + threadIdx.x;
    while (p < end) {
        const unsigned long long gp = (unsigned long long)(p + A.sample_offset);
        float y[NP];
        float b = a, tp = 0.f, fix = 0.f;
#pragma unroll
        for (int j = 0; j < NP; j++) y[j] = __fmul_rn(z[j], b);
        unsigned int blk = 0;
        for (;;) {
            bool hit = false;
#pragma unroll
            for (int j = 0; j < NP; j++) hit |= (y[j] > b);
            if (hit || !(tp <= A.max_t)) break;
            const float4 nn = normals4(philox4x32((unsigned int)gp, (unsigned int)(gp >> 32), (unsigned int)set, blk++, rk));
            float sum = 0.f;
            if (LCA) {
#pragma unroll
                for (int j = 0; j < NP; j++) sum = __fadd_rn(sum, y[j]);   // csum(), :62-70
            }
#pragma unroll
            for (int j = 0; j < NP; j++) {
                const float nj = (j == 0) ? nn.x : (j == 1) ? nn.y : (j == 2) ? nn.z : nn.w;
                if (LCA) {
                    const float red = __fadd_rn(-y[j], sum);
                    const float drift = __fmul_rn(__fsub_rn(__fsub_rn(v[j], __fmul_rn(g, y[j])), __fmul_rn(binh, red)), A.dt);
                    y[j] = fmaxf(0.0f, __fadd_rn(y[j], __fadd_rn(drift, __fmul_rn(sqrt_st, nj))));
                } else {
                    y[j] = __fadd_rn(y[j], __fadd_rn(vdt[j], __fmul_rn(sqrt_st, nj)));
                }
            }
            tp = __fadd_rn(tp, A.dt);
            fix += 1.0f;
            if (ANGLE) b = fmaf(fix, sdt, a);
        }
        my_steps += (unsigned int)fix;
        int choice = 0;                       // np.argmax: first index of the maximum
        float best = y[0];
#pragma unroll
        for (int j = 1; j < NP; j++)
            if (y[j] > best) { best = y[j]; choice = j; }
        const float rt = __fadd_rn(tp, t_nd);
        if (A.rts) {
            const size_t o = (size_t)p * A.n_sets + set;
            A.rts[o] = rt;
            A.choices[o] = choice;
        }
        if (A.hist) {
            const double rd = (double)rt;
            int bin = (int)(rt * A.hist_inv_step);   // within one bin of the truth; settled below
            if (bin > A.nbins - 1) bin = A.nbins - 1;
            if (bin < 0) bin = 0;
            const double e_lo = (bin == A.nbins - 1) ? A.hist_max_t : bin * A.hist_step;
            if (rd < e_lo && bin > 0) bin--;
            else if (bin < A.nbins - 1) {
                const double e_hi = (bin + 1 == A.nbins - 1) ? A.hist_max_t : (bin + 1) * A.hist_step;
                if (rd >= e_hi) bin++;
            }
            if (rd >= 0.0) atomicAdd(&s_hist[bin * NP + choice], 1u);
        }
        p = beg + (long long)atomicAdd(&s_next, 1);
    }
    __syncthreads();
    if (A.hist) {
        unsigned int* gh = A.hist + (size_t)set * nh;
        for (int i = threadIdx.x; i < nh; i += kSimBlock) {
            const unsigned int c = s_hist[i];
            if (c) atomicAdd(&gh[i], c);
        }
    }
    if (A.steps) {
        __shared__ unsigned long long s_steps[kSimBlock / 32];
        for (int off = 16; off > 0; off >>= 1) my_steps += __shfl_down_sync(0xffffffffu, my_steps, off);
        if ((threadIdx.x & 31) == 0) s_steps[threadIdx.x >> 5] = my_steps;
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned long long tot = 0;
            for (int i = 0; i < kSimBlock / 32; i++) tot += s_steps[i];
            atomicAdd(A.steps, tot);
        }
    }
}

template <int NP, bool LCA>
static cudaError_t launch_race(const Args& A, int angle, int blocks, cudaStream_t stream)
{
    const size_t smem = A.hist ? sizeof(unsigned int) * NP * (size_t)A.nbins : 0;
    if (angle) race_kernel<NP, LCA, true><<<blocks, kSimBlock, smem, stream>>>(A);
    else race_kernel<NP, LCA, false><<<blocks, kSimBlock, smem, stream>>>(A);
    return cudaGetLastError();
}

// ---- two-stage 4-choice model: ddm_flexbound_par2 (cddm_data_simulation.pyx:1667-1791) ------------------------------------
// A "high-dimension" walk (v_h, z_h) decides between the option pairs {0,1} and {2,3}; a "low-dimension" walk with the
// drift / start point of the chosen pair (v_l1, z_l1 or v_l2, z_l2) decides inside the pair.  Both walks start at time 0
// against the same bound b(t) (each with its own step index); rt = max(t_h, t_l) + t; choice = (y_h < 0 ? 0 : 2) +
// (sign(y_l) >= 0).  theta (canonical): v_h, v_l1, v_l2, a, z_h, z_l1, z_l2, t, b0, b1; BOUND 0 constant, 1 angle (b0),
// 2 Weibull (b0 = alpha, b1 = beta).  One Philox stream per particle, consumed in order by the two walks.
// KIND 0 is par2; KIND 1 / 2 / 3 are its siblings ddm_flexbound_seq2 / _mic2_adj / _tradeoff (theta gains an 11th column d),
// described where they branch off below.
template <int KIND, int BOUND>
__global__ void __launch_bounds__(kSimBlock) hier_kernel(const Args A)
{
    extern __shared__ unsigned int s_hist[];  // [nbins][4] when A.hist
    __shared__ int s_next;
    const int nh = A.hist ? A.nbins * 4 : 0;
    const float sqrt_st = __fmul_rn(sqrtf(A.dt), A.s);
    const RoundKeys rk(A);
    unsigned long long my_steps = 0;
    const long long item = blockIdx.x;
    const int set = (int)(item / A.items_per_set);
    const long long beg = (item - (long long)set * A.items_per_set) * A.chunk;
    const long long end = (beg + A.chunk < A.n_samples) ? beg + A.chunk : A.n_samples;
    for (int i = threadIdx.x; i < nh; i += kSimBlock) s_hist[i] = 0u;
    if (threadIdx.x == 0) s_next = kSimBlock;
    __syncthreads();

    const float* th = A.theta + (size_t)set * A.n_par;
    const float v_h = th[0], v_l1 = th[1], v_l2 = th[2], a = th[3], z_h = th[4], z_l1 = th[5], z_l2 = th[6], t_nd = th[7];
    const float sdt = (BOUND == 1 || BOUND == 3) ? (-sinf(th[8]) / cosf(th[8])) * A.dt : 0.f;
    const float w_alpha = th[8], w_inv_beta = (BOUND == 2) ? 1.0f / th[9] : 0.f;
    // conflict-gamma bound (graveyard/boundary_functions.py:66-97): max(a + scale * gamma_pdf(t; shape, scale_g) + t * slope, 0)
    const float g_scale = (BOUND == 3) ? th[9] : 0.f, g_shape_m1 = (BOUND == 3) ? th[11] - 1.f : 0.f;
    const float g_inv_sg = (BOUND == 3) ? 1.0f / th[12] : 0.f;
    const float g_lognorm = (BOUND == 3) ? -lgammaf(th[11]) + logf(g_inv_sg) : 0.f;
    auto bound = [&](float fix) -> float {
        if (BOUND == 1) return fmaf(fix, sdt, a);
        if (BOUND == 2) {
            const float r = (fix * A.dt) * w_inv_beta;
            const float pw = (r > 0.f) ? exp2f(w_alpha * __log2f(r)) : 0.f;
            return a * __expf(-pw);
        }
        if (BOUND == 3) {
            const float x = (fix * A.dt) * g_inv_sg;
            float pdf;
            if (x > 0.f) pdf = __expf(fmaf(g_shape_m1, __logf(x), g_lognorm - x));
            else pdf = (g_shape_m1 > 0.f) ? 0.f : (g_shape_m1 == 0.f ? g_inv_sg : __int_as_float(0x7f800000));
            return fmaxf(fmaf(g_scale, pdf, fmaf(fix, sdt, a)), 0.f);
        }
        return a;
    };
    const float b0 = bound(0.f);

    // Counter word 3 = block index | stream << 28: separate streams for the high-dimension walk (0), the low-dimension
    // walk (1) and the uniforms (2), so that the high-dimension path can be regenerated while the low one runs.
    struct Stream {
        unsigned int blk;
        int have;
        float4 nn;
    };
    const float d_damp = (KIND >= 2) ? th[10] : 0.f;
    const float vdt_h = __fmul_rn(v_h, A.dt);

    long long p = beg + threadIdx.x;
    while (p < end) {
        const unsigned long long gp = (unsigned long long)(p + A.sample_offset);
        float rt;
        int choice;
        if (KIND == 0) {
            Stream S = {0u, 0, make_float4(0.f, 0.f, 0.f, 0.f)};
            auto normal = [&]() -> float {
                if (S.have == 0) {
                    S.nn = normals4(philox4x32((unsigned int)gp, (unsigned int)(gp >> 32), (unsigned int)set, S.blk++, rk));
                    S.have = 4;
                }
                S.have--;
                return S.have == 3 ? S.nn.x : S.have == 2 ? S.nn.y : S.have == 1 ? S.nn.z : S.nn.w;
            };
            auto walk = [&](float z, float v, float* t_out) -> float {
                float y = __fadd_rn(-b0, __fmul_rn(__fmul_rn(z, 2.f), b0));
                float b = b0, tp = 0.f, fix = 0.f;
                const float vdt = __fmul_rn(v, A.dt);
                while (y >= -b && y <= b && tp <= A.max_t) {
                    y = __fadd_rn(y, __fadd_rn(vdt, __fmul_rn(sqrt_st, normal())));
                    tp = __fadd_rn(tp, A.dt);
                    fix += 1.0f;
                    if (BOUND != 0) b = bound(fix);
                }
                my_steps += (unsigned int)fix;
                *t_out = tp;
                return y;
            };
            float t_h, t_l;
            const float y_h = walk(z_h, v_h, &t_h);
            const bool lower = y_h < 0.f;                       // sign(y_h) < 0  (:1747)
            const float y_l = walk(lower ? z_l1 : z_l2, lower ? v_l1 : v_l2, &t_l);
            choice = (lower ? 0 : 2) + ((y_l >= 0.f) ? 1 : 0);
            rt = __fadd_rn(fmaxf(t_h, t_l), t_nd);
        } else {
            auto draw = [&](Stream& S, unsigned long long q, unsigned int stream) -> float {
                if (S.have == 0) {
                    S.nn = normals4(philox4x32((unsigned int)q, (unsigned int)(q >> 32), (unsigned int)set,
                                               (S.blk++) | (stream << 28), rk));
                    S.have = 4;
                }
                S.have--;
                return S.have == 3 ? S.nn.x : S.have == 2 ? S.nn.y : S.have == 1 ? S.nn.z : S.nn.w;
            };
            auto uniforms = [&](unsigned long long q) -> uint4 {
                return philox4x32((unsigned int)q, (unsigned int)(q >> 32), (unsigned int)set, 2u << 28, rk);
            };
            if (KIND == 1) {
                // seq2 (:1523-1661): the low-dimension walk continues on the clock and bound index of the high one.  On the
                // time-out and collapsed-bound branches the reference leaves y_l as the previous particle left it, so the
                // last increment reads the predecessor's final y_l: replay predecessors (by particle index) until one of
                // them ran its own low-dimension walk.  Particle 0 has none: y_l = 0 there.
                auto particle = [&](unsigned long long q, float& rt_q, int& base_q, bool& yl_nonneg, bool count) -> bool {
                    Stream H = {0u, 0, make_float4(0.f, 0.f, 0.f, 0.f)};
                    float y = __fadd_rn(-b0, __fmul_rn(__fmul_rn(z_h, 2.f), b0));
                    float b = b0, tp = 0.f, fix = 0.f;
                    while (y >= -b && y <= b && tp <= A.max_t) {
                        y = __fadd_rn(y, __fadd_rn(vdt_h, __fmul_rn(sqrt_st, draw(H, q, 0u))));
                        tp = __fadd_rn(tp, A.dt);
                        fix += 1.0f;
                        if (BOUND != 0) b = bound(fix);
                    }
                    bool stale;
                    float y_l = 0.f;
                    if (tp >= A.max_t) {
                        base_q = (u01(uniforms(q).x) > 0.5f) ? 1 : 0;
                        stale = true;
                    } else {
                        const bool lower = y < 0.f;
                        const float z_l = lower ? z_l1 : z_l2;
                        base_q = lower ? 0 : 2;
                        if (b <= 0.f) {
                            if (u01(uniforms(q).x) < z_l) base_q += 1;
                            stale = true;
                        } else {
                            stale = false;
                            y_l = __fadd_rn(-b, __fmul_rn(__fmul_rn(z_l, 2.f), b));
                            const float vdt = __fmul_rn(lower ? v_l1 : v_l2, A.dt);
                            Stream Lw = {0u, 0, make_float4(0.f, 0.f, 0.f, 0.f)};
                            while (y_l >= -b && y_l <= b && tp <= A.max_t) {
                                y_l = __fadd_rn(y_l, __fadd_rn(vdt, __fmul_rn(sqrt_st, draw(Lw, q, 1u))));
                                tp = __fadd_rn(tp, A.dt);
                                fix += 1.0f;
                                if (BOUND != 0) b = bound(fix);
                            }
                        }
                    }
                    if (count) my_steps += (unsigned int)fix;
                    rt_q = __fadd_rn(tp, t_nd);
                    yl_nonneg = y_l >= 0.f;
                    return stale;
                };
                bool nonneg;
                bool stale = particle(gp, rt, choice, nonneg, true);
                unsigned long long q = gp;
                for (int depth = 0; stale; depth++) {
                    if (q == 0ull || depth == 256) { nonneg = true; break; }
                    q--;
                    float rt_q;
                    int base_q;
                    stale = particle(q, rt_q, base_q, nonneg, false);
                }
                choice += nonneg ? 1 : 0;
            } else {
                // mic2_adj (:1797-1960) / tradeoff (:1966-2137): both walks start at time 0; while the high-dimension walk
                // is still running, the low one is slowed by the high walker's normalised position ("bias trace").  The
                // trace is not stored: the high walk is regenerated from its own stream next to the low walk.  Entry n_h
                // of the reference's trace is left over from the previous particle; it is taken as "already decided".
                Stream H = {0u, 0, make_float4(0.f, 0.f, 0.f, 0.f)};
                float y = __fadd_rn(-b0, __fmul_rn(__fmul_rn(z_h, 2.f), b0));
                float b = b0, t_h = 0.f, n_h = 0.f;
                while (y >= -b && y <= b && t_h <= A.max_t) {
                    y = __fadd_rn(y, __fadd_rn(vdt_h, __fmul_rn(sqrt_st, draw(H, gp, 0u))));
                    t_h = __fadd_rn(t_h, A.dt);
                    n_h += 1.0f;
                    if (BOUND != 0) b = bound(n_h);
                }
                const uint4 uu = uniforms(gp);
                const bool upper = u01(uu.x) <= __fdiv_rn(__fadd_rn(y, b), __fmul_rn(2.f, b));
                const float v_l = upper ? v_l2 : v_l1;
                const float vdt_l = __fmul_rn(v_l, A.dt);
                float y_l = __fadd_rn(-b0, __fmul_rn(__fmul_rn(upper ? z_l2 : z_l1, 2.f), b0));
                Stream H2 = {0u, 0, make_float4(0.f, 0.f, 0.f, 0.f)}, Lw = H2;
                float y2 = __fadd_rn(-b0, __fmul_rn(__fmul_rn(z_h, 2.f), b0));
                float t_l = 0.f, fix = 0.f;
                b = b0;
                const float inv_2md = __fdiv_rn(1.f, __fsub_rn(2.f, d_damp));
                while (y_l >= -b && y_l <= b && t_l <= A.max_t) {
                    float tr = 1.f;
                    if (fix < n_h) {
                        y2 = __fadd_rn(y2, __fadd_rn(vdt_h, __fmul_rn(sqrt_st, draw(H2, gp, 0u))));
                        tr = __fdiv_rn(__fadd_rn(y2, b), __fmul_rn(2.f, b));
                        if (!upper) tr = __fsub_rn(1.f, tr);
                    }
                    const bool before = (tr < 1.f) && (tr > 0.f);
                    const float g = draw(Lw, gp, 1u);
                    if (KIND == 2) {
                        y_l = __fadd_rn(y_l, before ? __fmul_rn(__fmul_rn(__fmul_rn(v_l, tr), d_damp), A.dt) : vdt_l);
                        y_l = __fadd_rn(y_l, __fmul_rn(sqrt_st, g));
                    } else {
                        const float pos = before ? __fmul_rn(__fadd_rn(1.f, __fmul_rn(d_damp, __fsub_rn(tr, 1.f))), inv_2md) : 1.f;
                        y_l = __fadd_rn(y_l, __fmul_rn(pos, vdt_l));
                        y_l = __fadd_rn(y_l, __fmul_rn(__fmul_rn(pos, sqrt_st), g));
                    }
                    t_l = __fadd_rn(t_l, A.dt);
                    fix += 1.0f;
                    if (BOUND != 0) b = bound(fix);
                }
                my_steps += (unsigned long long)n_h + (unsigned long long)fix;
                choice = (upper ? 2 : 0) + ((u01(uu.y) <= __fdiv_rn(__fadd_rn(y_l, b), __fmul_rn(2.f, b))) ? 1 : 0);
                rt = __fadd_rn(fmaxf(t_h, t_l), t_nd);
            }
        }
        if (A.rts) {
            const size_t o = (size_t)p * A.n_sets + set;
            A.rts[o] = rt;
            A.choices[o] = choice;
        }
        if (A.hist) {
            const double rd = (double)rt;
            int bin = (int)(rt * A.hist_inv_step);   // within one bin of the truth; settled below
            if (bin > A.nbins - 1) bin = A.nbins - 1;
            if (bin < 0) bin = 0;
            const double e_lo = (bin == A.nbins - 1) ? A.hist_max_t : bin * A.hist_step;
            if (rd < e_lo && bin > 0) bin--;
            else if (bin < A.nbins - 1) {
                const double e_hi = (bin + 1 == A.nbins - 1) ? A.hist_max_t : (bin + 1) * A.hist_step;
                if (rd >= e_hi) bin++;
            }
            if (rd >= 0.0) atomicAdd(&s_hist[bin * 4 + choice], 1u);
        }
        p = beg + (long long)atomicAdd(&s_next, 1);
    }
    __syncthreads();
    if (A.hist) {
        unsigned int* gh = A.hist + (size_t)set * nh;
        for (int i = threadIdx.x; i < nh; i += kSimBlock) {
            const unsigned int c = s_hist[i];
            if (c) atomicAdd(&gh[i], c);
        }
    }
    if (A.steps) {
        __shared__ unsigned long long s_steps[kSimBlock / 32];
        for (int off = 16; off > 0; off >>= 1) my_steps += __shfl_down_sync(0xffffffffu, my_steps, off);
        if ((threadIdx.x & 31) == 0) s_steps[threadIdx.x >> 5] = my_steps;
        __syncthreads();
        if (threadIdx.x == 0) {
            unsigned long long tot = 0;
            for (int i = 0; i < kSimBlock / 32; i++) tot += s_steps[i];
            atomicAdd(A.steps, tot);
        }
    }
}

// Particles per work item (one CTA each): 4,096 for large runs; for small ones fewer, so that a call with a few thousand
// particles still spreads over the machine instead of running on a single SM (results do not depend on it: the random
// streams are keyed by particle index).
static long long pick_chunk(long long n_samples, long long n_sets)
{
    long long chunk = 4096;
    const long long want_items = (long long)api_sm_count() * 4;
    // plenty of work: longer items, so that the tail of an item (lanes waiting for its last walks) weighs less
    while (chunk < 32768 && n_sets * ((n_samples + 2 * chunk - 1) / (2 * chunk)) >= 16 * want_items) chunk *= 2;
    if (n_sets * ((n_samples + chunk - 1) / chunk) < want_items) {
        chunk = (n_samples * n_sets + want_items - 1) / want_items;
        if (chunk < kSimBlock) chunk = kSimBlock;
        if (chunk > 4096) chunk = 4096;
    }
    if (chunk > n_samples) chunk = n_samples;
    return chunk;
}

template <int MODEL>
static cudaError_t launch(const Args& A, int blocks, cudaStream_t stream)
{
    const size_t smem = A.hist ? sizeof(unsigned int) * 2 * (size_t)A.nbins : 0;
    Args P = A;
    P.traj = nullptr;
    if (A.traj && A.sample_offset == 0) {
        // the trajectory of (set 0, particle 0): the same walk again, one particle, path written out.  One thread
        // stepping through a whole walk takes as long as a small production launch, so it runs beside it on a stream
        // of its own (forked from and joined back into the caller's stream with events).
        static cudaStream_t side[16] = {};
        static cudaEvent_t fork_ev[16] = {}, join_ev[16] = {};
        int dev = 0;
        cudaGetDevice(&dev);
        dev &= 15;
        if (!side[dev]) {
            cudaStreamCreateWithFlags(&side[dev], cudaStreamNonBlocking);
            cudaEventCreateWithFlags(&fork_ev[dev], cudaEventDisableTiming);
            cudaEventCreateWithFlags(&join_ev[dev], cudaEventDisableTiming);
        }
        Args T = A;
        T.n_sets = 1; T.n_samples = 1; T.chunk = 1; T.items_per_set = 1;
        T.rts = nullptr; T.choices = nullptr; T.hist = nullptr; T.nbins = 0; T.steps = nullptr;
        cudaEventRecord(fork_ev[dev], stream);            // inputs (theta, the -999 prefill) are ready at this point
        cudaStreamWaitEvent(side[dev], fork_ev[dev], 0);
        sim_kernel<MODEL, true><<<1, kSimBlock, 0, side[dev]>>>(T);
        cudaEventRecord(join_ev[dev], side[dev]);
        sim_kernel<MODEL, false><<<blocks, kSimBlock, smem, stream>>>(P);
        cudaStreamWaitEvent(stream, join_ev[dev], 0);
        return cudaGetLastError();
    }
    sim_kernel<MODEL, false><<<blocks, kSimBlock, smem, stream>>>(P);
    return cudaGetLastError();
}

static int n_params_of(int model)
{
    switch (model) {
        case WFPT_SIM_DDM: case WFPT_SIM_DDM_FLEXBOUND: case WFPT_SIM_DDM_COV: return 4;
        case WFPT_SIM_GAMMA_DRIFT: return 7;
        case WFPT_SIM_GAMMA_DRIFT_ANGLE: return 8;
        case WFPT_SIM_DS_CONFLICT: return 10;
        case WFPT_SIM_DS_CONFLICT_ANGLE: return 11;
        case WFPT_SIM_ANGLE: case WFPT_SIM_DDM_SDV: case WFPT_SIM_ORNSTEIN: case WFPT_SIM_LEVY: return 5;
        case WFPT_SIM_WEIBULL: return 6;
        case WFPT_SIM_FULL_DDM_BASE: case WFPT_SIM_FULL_DDM: return 7;
        default: return -1;
    }
}

}  // namespace sim
}  // namespace wfpt

using namespace wfpt;

extern "C" int wfpt_sim_n_params(int model) { return sim::n_params_of(model); }

extern "C" int wfpt_sim_run(int model, const float* d_theta, int n_par, int n_sets, int64_t n_samples, int64_t sample_offset,
                            float delta_t, float max_t, float s, uint64_t seed, float* d_rts, int32_t* d_choices,
                            uint32_t* d_hist, int nbins, float hist_max_t, uint64_t* d_steps, float* d_traj, int traj_len,
                            void* stream)
{
    return wfpt_sim_run_sharded(model, d_theta, n_par, n_sets, 0, 1, n_samples, sample_offset, delta_t, max_t, s, seed, d_rts,
                                d_choices, d_hist, nbins, hist_max_t, d_steps, d_traj, traj_len, stream);
}

extern "C" int wfpt_sim_run_sharded(int model, const float* d_theta, int n_par, int n_sets, int set_offset, int set_stride,
                                    int64_t n_samples, int64_t sample_offset, float delta_t, float max_t, float s,
                                    uint64_t seed, float* d_rts, int32_t* d_choices, uint32_t* d_hist, int nbins,
                                    float hist_max_t, uint64_t* d_steps, float* d_traj, int traj_len, void* stream)
{
    if (set_offset < 0 || set_stride < 1) return api_fail("wfpt_sim_run_sharded: set_offset must be >= 0 and set_stride >= 1");
    const int want = sim::n_params_of(model);
    if (want < 0) { char m[96]; snprintf(m, sizeof m, "wfpt_sim_run: unknown model id %d", model); return api_fail(m); }
    if (n_par != want) { char m[128]; snprintf(m, sizeof m, "wfpt_sim_run: model %d takes %d parameters per set, got %d", model, want, n_par); return api_fail(m); }
    if (n_sets < 0 || n_samples < 0) return api_fail("wfpt_sim_run: negative size");
    if (n_sets == 0 || n_samples == 0) return 0;
    if (!d_theta) return api_fail("wfpt_sim_run: null theta");
    if ((d_rts == nullptr) != (d_choices == nullptr)) return api_fail("wfpt_sim_run: rts and choices must be given together");
    if (!d_rts && !d_hist) return api_fail("wfpt_sim_run: nothing to write (no rts/choices and no histogram)");
    if (!(delta_t > 0.f) || !(max_t > 0.f)) return api_fail("wfpt_sim_run: delta_t and max_t must be positive");
    if (d_hist && (nbins < 2 || nbins > sim::kMaxBins)) return api_fail("wfpt_sim_run: nbins must be in [2, 4096]");
    if (d_hist && !(hist_max_t > 0.f)) return api_fail("wfpt_sim_run: hist_max_t must be positive");

    sim::Args A;
    A.theta = d_theta; A.n_par = n_par; A.n_sets = n_sets; A.n_samples = n_samples; A.sample_offset = sample_offset;
    A.set_offset = set_offset; A.set_stride = set_stride;
    A.out_stride_samples = n_samples;
    A.dt = delta_t; A.max_t = max_t; A.s = s; A.set_seed(seed); A.rts = d_rts; A.choices = d_choices;
    A.hist = d_hist; A.nbins = d_hist ? nbins : 0;
    A.hist_max_t = (double)hist_max_t;
    A.hist_step = d_hist ? (double)hist_max_t / (double)(nbins - 1) : 1.0;
    A.hist_inv_step = (float)(1.0 / A.hist_step);
    A.steps = (unsigned long long*)d_steps;
    A.traj = d_traj; A.traj_len = traj_len;

    int dev = 0, sms = 148;
    if (cudaGetDevice(&dev) == cudaSuccess) cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    // work items: (parameter set, chunk of its particles), one CTA each.  Chunks of at most 4,096 particles keep the
    // hardware block scheduler busy balancing cheap and expensive parameter sets, while inside a CTA the shared
    // particle counter keeps all lanes busy until the chunk is done.
    const long long chunk = sim::pick_chunk(n_samples, n_sets);
    A.chunk = chunk;
    A.items_per_set = (n_samples + chunk - 1) / chunk;
    const long long n_items = A.items_per_set * n_sets;
    if (n_items > 2147483647LL) return api_fail("wfpt_sim_run: too many work items for one launch; split the call");
    const int blocks = (int)n_items;
    (void)sms;

    cudaError_t e;
    cudaStream_t st = (cudaStream_t)stream;
    switch (model) {
        case WFPT_SIM_DDM: e = sim::launch<WFPT_SIM_DDM>(A, blocks, st); break;
        case WFPT_SIM_FULL_DDM_BASE: e = sim::launch<WFPT_SIM_FULL_DDM_BASE>(A, blocks, st); break;
        case WFPT_SIM_DDM_FLEXBOUND: e = sim::launch<WFPT_SIM_DDM_FLEXBOUND>(A, blocks, st); break;
        case WFPT_SIM_ANGLE: e = sim::launch<WFPT_SIM_ANGLE>(A, blocks, st); break;
        case WFPT_SIM_WEIBULL: e = sim::launch<WFPT_SIM_WEIBULL>(A, blocks, st); break;
        case WFPT_SIM_FULL_DDM: e = sim::launch<WFPT_SIM_FULL_DDM>(A, blocks, st); break;
        case WFPT_SIM_DDM_SDV: e = sim::launch<WFPT_SIM_DDM_SDV>(A, blocks, st); break;
        case WFPT_SIM_ORNSTEIN: e = sim::launch<WFPT_SIM_ORNSTEIN>(A, blocks, st); break;
        case WFPT_SIM_GAMMA_DRIFT: e = sim::launch<WFPT_SIM_GAMMA_DRIFT>(A, blocks, st); break;
        case WFPT_SIM_GAMMA_DRIFT_ANGLE: e = sim::launch<WFPT_SIM_GAMMA_DRIFT_ANGLE>(A, blocks, st); break;
        case WFPT_SIM_DS_CONFLICT: e = sim::launch<WFPT_SIM_DS_CONFLICT>(A, blocks, st); break;
        case WFPT_SIM_DS_CONFLICT_ANGLE: e = sim::launch<WFPT_SIM_DS_CONFLICT_ANGLE>(A, blocks, st); break;
        case WFPT_SIM_DDM_COV: e = sim::launch<WFPT_SIM_DDM_COV>(A, blocks, st); break;
        default: e = sim::launch<WFPT_SIM_LEVY>(A, blocks, st); break;
    }
    if (e != cudaSuccess) return api_fail(cudaGetErrorString(e));
    count_launch(1);
    return 0;
}

#define SCK(call)                                                      \
    do {                                                               \
        cudaError_t e__ = (call);                                      \
        if (e__ != cudaSuccess) { rc = api_fail(cudaGetErrorString(e__)); goto done; } \
    } while (0)

#define SCK_RC(call)                        \
    do {                                    \
        if ((call) != 0) { rc = 1; goto done; } \
    } while (0)

extern "C" int wfpt_sim_run_host(int model, const float* h_theta, int n_par, int n_sets, int64_t n_samples, int64_t sample_offset,
                                 float delta_t, float max_t, float s, uint64_t seed, float* h_rts, int32_t* h_choices,
                                 uint32_t* h_hist, int nbins, float hist_max_t, uint64_t* h_steps, float* h_traj, int traj_len)
{
    wfpt::HostCall guard_;
    if (n_sets <= 0 || n_samples <= 0) return wfpt_sim_run(model, nullptr, n_par, n_sets, n_samples, sample_offset, delta_t, max_t, s, seed,
                                                           nullptr, nullptr, nullptr, nbins, hist_max_t, nullptr, nullptr, 0, nullptr);
    int rc = 0;
    float *d_theta = nullptr, *d_rts = nullptr, *d_traj = nullptr;
    int32_t* d_ch = nullptr;
    uint32_t* d_hist = nullptr;
    uint64_t* d_steps = nullptr;
    const size_t n_out = (size_t)n_samples * (size_t)n_sets;
    const size_t hist_n = (size_t)n_sets * (size_t)(nbins > 0 ? nbins : 0) * 2;
    SCK_RC(api_sim_scratch(0, sizeof(float) * (size_t)n_sets * n_par, (void**)&d_theta));
    SCK(cudaMemcpy(d_theta, h_theta, sizeof(float) * (size_t)n_sets * n_par, cudaMemcpyHostToDevice));
    if (h_rts) { SCK_RC(api_sim_scratch(1, sizeof(float) * n_out, (void**)&d_rts)); SCK_RC(api_sim_scratch(2, sizeof(int32_t) * n_out, (void**)&d_ch)); }
    if (h_hist) { SCK_RC(api_sim_scratch(3, sizeof(uint32_t) * hist_n, (void**)&d_hist)); SCK(cudaMemset(d_hist, 0, sizeof(uint32_t) * hist_n)); }
    if (h_steps) { SCK_RC(api_sim_scratch(4, sizeof(uint64_t), (void**)&d_steps)); SCK(cudaMemset(d_steps, 0, sizeof(uint64_t))); }
    if (h_traj && traj_len > 0) {
        SCK_RC(api_sim_scratch(5, sizeof(float) * (size_t)traj_len, (void**)&d_traj));
        // the reference pre-fills the trajectory with -999 (cddm_data_simulation.pyx:318-319)
        SCK(cudaMemcpy(d_traj, h_traj, sizeof(float) * (size_t)traj_len, cudaMemcpyHostToDevice));
    }
    rc = wfpt_sim_run(model, d_theta, n_par, n_sets, n_samples, sample_offset, delta_t, max_t, s, seed, d_rts, d_ch, d_hist, nbins,
                      hist_max_t, d_steps, d_traj, traj_len, nullptr);
    if (rc) goto done;
    SCK(cudaDeviceSynchronize());
    if (h_rts) {
        SCK(cudaMemcpy(h_rts, d_rts, sizeof(float) * n_out, cudaMemcpyDeviceToHost));
        SCK(cudaMemcpy(h_choices, d_ch, sizeof(int32_t) * n_out, cudaMemcpyDeviceToHost));
    }
    if (h_hist) SCK(cudaMemcpy(h_hist, d_hist, sizeof(uint32_t) * hist_n, cudaMemcpyDeviceToHost));
    if (h_steps) SCK(cudaMemcpy(h_steps, d_steps, sizeof(uint64_t), cudaMemcpyDeviceToHost));
    if (d_traj) SCK(cudaMemcpy(h_traj, d_traj, sizeof(float) * (size_t)traj_len, cudaMemcpyDeviceToHost));
done:
    return rc;
}


extern "C" int wfpt_sim_race_run_host(int lca, int n_particles, int angle, const float* h_theta, int n_sets, int64_t n_samples,
                                      int64_t sample_offset, float delta_t, float max_t, float s, uint64_t seed, float* h_rts,
                                      int32_t* h_choices, uint32_t* h_hist, int nbins, float hist_max_t, uint64_t* h_steps)
{
    wfpt::HostCall guard_;
    if (n_particles < 2 || n_particles > 4) return api_fail("wfpt_sim_race_run_host: n_particles must be 2, 3 or 4");
    if (n_sets < 0 || n_samples < 0) return api_fail("wfpt_sim_race_run_host: negative size");
    if (n_sets == 0 || n_samples == 0) return 0;
    if (!h_theta) return api_fail("wfpt_sim_race_run_host: null theta");
    if ((h_rts == nullptr) != (h_choices == nullptr)) return api_fail("wfpt_sim_race_run_host: rts and choices must be given together");
    if (!h_rts && !h_hist) return api_fail("wfpt_sim_race_run_host: nothing to write");
    if (!(delta_t > 0.f) || !(max_t > 0.f)) return api_fail("wfpt_sim_race_run_host: delta_t and max_t must be positive");
    if (h_hist && (nbins < 2 || nbins * n_particles > 2 * sim::kMaxBins)) return api_fail("wfpt_sim_race_run_host: nbins out of range");
    const int n_par = 2 * n_particles + 5;
    int rc = 0;
    float *d_theta = nullptr, *d_rts = nullptr;
    int32_t* d_ch = nullptr;
    uint32_t* d_hist = nullptr;
    uint64_t* d_steps = nullptr;
    const size_t n_out = (size_t)n_samples * (size_t)n_sets;
    const size_t hist_n = (size_t)n_sets * (size_t)(nbins > 0 ? nbins : 0) * n_particles;
    sim::Args A;
    cudaError_t e = cudaSuccess;
    SCK_RC(api_sim_scratch(0, sizeof(float) * (size_t)n_sets * n_par, (void**)&d_theta));
    SCK(cudaMemcpy(d_theta, h_theta, sizeof(float) * (size_t)n_sets * n_par, cudaMemcpyHostToDevice));
    if (h_rts) { SCK_RC(api_sim_scratch(1, sizeof(float) * n_out, (void**)&d_rts)); SCK_RC(api_sim_scratch(2, sizeof(int32_t) * n_out, (void**)&d_ch)); }
    if (h_hist) { SCK_RC(api_sim_scratch(3, sizeof(uint32_t) * hist_n, (void**)&d_hist)); SCK(cudaMemset(d_hist, 0, sizeof(uint32_t) * hist_n)); }
    if (h_steps) { SCK_RC(api_sim_scratch(4, sizeof(uint64_t), (void**)&d_steps)); SCK(cudaMemset(d_steps, 0, sizeof(uint64_t))); }
    A.theta = d_theta; A.n_par = n_par; A.n_sets = n_sets; A.n_samples = n_samples; A.sample_offset = sample_offset;
    A.out_stride_samples = n_samples; A.dt = delta_t; A.max_t = max_t; A.s = s; A.set_seed(seed); A.rts = d_rts; A.choices = d_ch;
    A.hist = d_hist; A.nbins = d_hist ? nbins : 0; A.hist_max_t = (double)hist_max_t;
    A.hist_step = d_hist ? (double)hist_max_t / (double)(nbins - 1) : 1.0;
    A.hist_inv_step = (float)(1.0 / A.hist_step);
    A.steps = (unsigned long long*)d_steps; A.traj = nullptr; A.traj_len = 0;
    A.chunk = sim::pick_chunk(n_samples, n_sets);
    A.items_per_set = (n_samples + A.chunk - 1) / A.chunk;
    {
        const long long n_items = A.items_per_set * n_sets;
        if (n_items > 2147483647LL) { rc = api_fail("wfpt_sim_race_run_host: too many work items; split the call"); goto done; }
        const int blocks = (int)n_items;
        if (n_particles == 2) e = lca ? sim::launch_race<2, true>(A, angle, blocks, 0) : sim::launch_race<2, false>(A, angle, blocks, 0);
        else if (n_particles == 3) e = lca ? sim::launch_race<3, true>(A, angle, blocks, 0) : sim::launch_race<3, false>(A, angle, blocks, 0);
        else e = lca ? sim::launch_race<4, true>(A, angle, blocks, 0) : sim::launch_race<4, false>(A, angle, blocks, 0);
        SCK(e);
        count_launch(1);
    }
    SCK(cudaDeviceSynchronize());
    if (h_rts) {
        SCK(cudaMemcpy(h_rts, d_rts, sizeof(float) * n_out, cudaMemcpyDeviceToHost));
        SCK(cudaMemcpy(h_choices, d_ch, sizeof(int32_t) * n_out, cudaMemcpyDeviceToHost));
    }
    if (h_hist) SCK(cudaMemcpy(h_hist, d_hist, sizeof(uint32_t) * hist_n, cudaMemcpyDeviceToHost));
    if (h_steps) SCK(cudaMemcpy(h_steps, d_steps, sizeof(uint64_t), cudaMemcpyDeviceToHost));
done:
    return rc;
}


static int run_hier_host(const char* who, int kind, int bound_kind, const float* h_theta, int n_sets, int64_t n_samples,
                         int64_t sample_offset, float delta_t, float max_t, float s, uint64_t seed, float* h_rts,
                         int32_t* h_choices, uint32_t* h_hist, int nbins, float hist_max_t, uint64_t* h_steps)
{
    if (bound_kind < 0 || bound_kind > (kind == 0 ? 2 : 3)) return api_fail((std::string(who) + ": bound_kind out of range").c_str());
    if (n_sets < 0 || n_samples < 0) return api_fail((std::string(who) + ": negative size").c_str());
    if (n_sets == 0 || n_samples == 0) return 0;
    if (!h_theta) return api_fail((std::string(who) + ": null theta").c_str());
    if ((h_rts == nullptr) != (h_choices == nullptr)) return api_fail((std::string(who) + ": rts and choices must be given together").c_str());
    if (!h_rts && !h_hist) return api_fail((std::string(who) + ": nothing to write").c_str());
    if (!(delta_t > 0.f) || !(max_t > 0.f)) return api_fail((std::string(who) + ": delta_t and max_t must be positive").c_str());
    if (h_hist && (nbins < 2 || nbins * 4 > 2 * sim::kMaxBins)) return api_fail((std::string(who) + ": nbins out of range").c_str());
    const int n_par = kind == 0 ? 10 : 13;
    int rc = 0;
    float *d_theta = nullptr, *d_rts = nullptr;
    int32_t* d_ch = nullptr;
    uint32_t* d_hist = nullptr;
    uint64_t* d_steps = nullptr;
    const size_t n_out = (size_t)n_samples * (size_t)n_sets;
    const size_t hist_n = (size_t)n_sets * (size_t)(nbins > 0 ? nbins : 0) * 4;
    sim::Args A;
    SCK_RC(api_sim_scratch(0, sizeof(float) * (size_t)n_sets * n_par, (void**)&d_theta));
    SCK(cudaMemcpy(d_theta, h_theta, sizeof(float) * (size_t)n_sets * n_par, cudaMemcpyHostToDevice));
    if (h_rts) { SCK_RC(api_sim_scratch(1, sizeof(float) * n_out, (void**)&d_rts)); SCK_RC(api_sim_scratch(2, sizeof(int32_t) * n_out, (void**)&d_ch)); }
    if (h_hist) { SCK_RC(api_sim_scratch(3, sizeof(uint32_t) * hist_n, (void**)&d_hist)); SCK(cudaMemset(d_hist, 0, sizeof(uint32_t) * hist_n)); }
    if (h_steps) { SCK_RC(api_sim_scratch(4, sizeof(uint64_t), (void**)&d_steps)); SCK(cudaMemset(d_steps, 0, sizeof(uint64_t))); }
    A.theta = d_theta; A.n_par = n_par; A.n_sets = n_sets; A.n_samples = n_samples; A.sample_offset = sample_offset;
    A.out_stride_samples = n_samples; A.dt = delta_t; A.max_t = max_t; A.s = s; A.set_seed(seed); A.rts = d_rts; A.choices = d_ch;
    A.hist = d_hist; A.nbins = d_hist ? nbins : 0; A.hist_max_t = (double)hist_max_t;
    A.hist_step = d_hist ? (double)hist_max_t / (double)(nbins - 1) : 1.0;
    A.hist_inv_step = (float)(1.0 / A.hist_step);
    A.steps = (unsigned long long*)d_steps; A.traj = nullptr; A.traj_len = 0;
    A.chunk = sim::pick_chunk(n_samples, n_sets);
    A.items_per_set = (n_samples + A.chunk - 1) / A.chunk;
    {
        const long long n_items = A.items_per_set * n_sets;
        if (n_items > 2147483647LL) { rc = api_fail((std::string(who) + ": too many work items; split the call").c_str()); goto done; }
        const size_t smem = A.hist ? sizeof(unsigned int) * 4 * (size_t)A.nbins : 0;
#define HIER_LAUNCH(K)                                                                           \
    do {                                                                                         \
        if (bound_kind == 0) sim::hier_kernel<K, 0><<<(int)n_items, sim::kSimBlock, smem>>>(A);  \
        else if (bound_kind == 1) sim::hier_kernel<K, 1><<<(int)n_items, sim::kSimBlock, smem>>>(A); \
        else if (bound_kind == 2) sim::hier_kernel<K, 2><<<(int)n_items, sim::kSimBlock, smem>>>(A); \
        else sim::hier_kernel<K, 3><<<(int)n_items, sim::kSimBlock, smem>>>(A);                  \
    } while (0)
        if (kind == 0) HIER_LAUNCH(0);
        else if (kind == 1) HIER_LAUNCH(1);
        else if (kind == 2) HIER_LAUNCH(2);
        else HIER_LAUNCH(3);
#undef HIER_LAUNCH
        SCK(cudaGetLastError());
        count_launch(1);
    }
    SCK(cudaDeviceSynchronize());
    if (h_rts) {
        SCK(cudaMemcpy(h_rts, d_rts, sizeof(float) * n_out, cudaMemcpyDeviceToHost));
        SCK(cudaMemcpy(h_choices, d_ch, sizeof(int32_t) * n_out, cudaMemcpyDeviceToHost));
    }
    if (h_hist) SCK(cudaMemcpy(h_hist, d_hist, sizeof(uint32_t) * hist_n, cudaMemcpyDeviceToHost));
    if (h_steps) SCK(cudaMemcpy(h_steps, d_steps, sizeof(uint64_t), cudaMemcpyDeviceToHost));
done:
    return rc;
}

extern "C" int wfpt_sim_par2_run_host(int bound_kind, const float* h_theta, int n_sets, int64_t n_samples, int64_t sample_offset,
                                      float delta_t, float max_t, float s, uint64_t seed, float* h_rts, int32_t* h_choices,
                                      uint32_t* h_hist, int nbins, float hist_max_t, uint64_t* h_steps)
{
    wfpt::HostCall guard_;
    return run_hier_host("wfpt_sim_par2_run_host", 0, bound_kind, h_theta, n_sets, n_samples, sample_offset, delta_t, max_t, s,
                         seed, h_rts, h_choices, h_hist, nbins, hist_max_t, h_steps);
}

extern "C" int wfpt_sim_hier_run_host(int kind, int bound_kind, const float* h_theta, int n_sets, int64_t n_samples,
                                      int64_t sample_offset, float delta_t, float max_t, float s, uint64_t seed, float* h_rts,
                                      int32_t* h_choices, uint32_t* h_hist, int nbins, float hist_max_t, uint64_t* h_steps)
{
    wfpt::HostCall guard_;
    if (kind < 1 || kind > 3) return api_fail("wfpt_sim_hier_run_host: kind must be 1 (seq2), 2 (mic2_adj) or 3 (tradeoff)");
    return run_hier_host("wfpt_sim_hier_run_host", kind, bound_kind, h_theta, n_sets, n_samples, sample_offset, delta_t, max_t, s,
                         seed, h_rts, h_choices, h_hist, nbins, hist_max_t, h_steps);
}
