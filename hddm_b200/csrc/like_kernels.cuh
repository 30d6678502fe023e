// Kernel skeletons of the likelihood path, generic over the density evaluator `Impl` (faithful or fast).
//
//   like_batch_kernel   wiener_like for P parameter sets on one RT array      (/root/reference/src/wfpt.pyx:54-76)
//   pdf_array_kernel    per-trial densities                                      (wfpt.pyx:32-48)
//   like_multi_kernel   per-trial parameters, +-999 coding, RL-DDM drift scaling (wfpt.pyx:357-387, :78-153)
//   finalize_kernel     second stage of the deterministic two-stage trial sum, for callers that do not pass arrival
//                       counters (like_batch_kernel otherwise does it itself: the slice that finishes last sums)
//
// Data layout in HBM: RTs are one contiguous float64[n]; parameter sets are float64[P][8] rows
// (v, sv, a, z, sz, t, st, p_outlier); outputs float64[P].  One CTA = (one parameter set, one slice of trials);
// its 128 threads walk the slice with coalesced 16-byte loads (two RTs per thread per step; one RT per thread and
// interleaved slices for small grids), reduce with warp shuffles, and either write the final sum (slices == 1) or a
// partial into scratch (slices > 1).
#pragma once
#include "common.cuh"

namespace wfpt {

// wfpt.pyx:70-74: p = p*(1-p_outlier) + w_outlier*p_outlier, each operation rounded on its own.
__device__ __forceinline__ double mix_outlier(double p, double one_minus_po, double wp)
{
    return __dadd_rn(__dmul_rn(p, one_minus_po), wp);
}

// sum_i log(p_i) accumulated as log(prod_i p_i): one multiply and a few integer instructions per trial instead of a
// 30-FMA log.  The running product is kept as a mantissa in [1, 2) plus an integer binary exponent.  Values outside
// the comfortable range (zero, negative, subnormal, huge, inf, NaN) take log() directly, which also keeps the
// reference's semantics for them (log(0) = -inf, log(negative) = NaN).  Error: n * 2^-53 absolute on the sum.
struct LogProduct {
    double m = 1.0;
    double slow = 0.0;
    long long e = 0;
    __device__ __forceinline__ void add(double p, int& zero)
    {
        const int hi = __double2hiint(p);
        if ((unsigned)(hi - 0x00200000) < 0x7fb00000u) {  // 2^-1020 <= p < 2^1022
            m *= p;
            const int hm = __double2hiint(m);
            e += (hm >> 20) - 1023;
            m = __hiloint2double((hm & 0x000fffff) | 0x3ff00000, __double2loint(m));
        } else {
            zero |= (p == 0);
            slow += log(p);
        }
    }
    __device__ __forceinline__ double value() const { return slow + fma((double)e, 0.6931471805599453, log(m)); }
};

// ADAPT: the launch uses adaptive Simpson (use_adaptive != 0).  Only then can a trial be deferred (refinement below
// the first level), so only that instantiation carries the parking lot, the round structure and the adaptive
// evaluators; the fixed-grid instantiation stays lean.
constexpr int kParkCheck = 3;   // rounds between two looks at the parking lot (capacity must exceed kParkCheck * 2 * kBlock)

// Resident CTAs per SM asked of the register allocator: the adaptive instantiation keeps its spill count at five (102
// registers: +8.5 % from the extra warps), the fixed-grid one loses more to spills than it gains (measured both ways).
template <class Impl, bool ADAPT, int MINB = (ADAPT ? 5 : 4)>
__global__ void __launch_bounds__(kBlock, MINB) like_batch_kernel(const double* __restrict__ x, int64_t n,
                                                          const double* __restrict__ params, int64_t P, Opts o,
                                                          double w_outlier, int slices, int64_t chunk,
                                                          double* __restrict__ out_sum, int* __restrict__ out_zero,
                                                          const int64_t* __restrict__ seg, int* __restrict__ done,
                                                          double* __restrict__ final_out, int one_each,
                                                          const __grid_constant__ InlineRow inl,
                                                          const __grid_constant__ typename Impl::Ctx ctx)
{
    const int64_t bid = blockIdx.x;
    const int64_t p = bid / slices;
    const int s = (int)(bid - p * slices);
    if (p >= P) return;
    if (seg) {
        // segmented form: every parameter set has its own RT array (one PyMC node each), sliced the same way
        const int64_t s0 = seg[2 * p];
        x += s0;
        n = seg[2 * p + 1] - s0;
        chunk = (n + slices - 1) / slices;
        if (!one_each) chunk = (chunk + 1) & ~(int64_t)1;   // even, so that 16-byte loads stay aligned across slices
        if (chunk < 2) chunk = 2;
    }

    // (a single parameter set given by a host caller rides in the kernel arguments: no memory read before the first trial)
    const double* row = inl.use ? inl.v : params + 8 * p;
    Params prm{row[0], row[1], row[2], row[3], row[4], row[5], row[6], row[7]};
    const double po = prm.p_outlier;

    // wfpt.pyx:63-64: p_outlier outside [0,1] -> -inf for the whole set
    if (!((po >= 0) & (po <= 1))) {
        if (threadIdx.x == 0) {
            if (out_zero) { out_sum[bid] = 0.0; out_zero[bid] = 1; }
            else out_sum[bid] = -CUDART_INF;
            if (out_zero && done) final_out[p] = -CUDART_INF;   // every slice of the set comes through here: no counting
        }
        return;
    }
    const double wp = __dmul_rn(w_outlier, po);
    const double omp = __dsub_rn(1.0, po);
    typename Impl::Set set;
    Impl::prepare(set, prm, o, ctx);

    const int64_t beg = (int64_t)s * chunk;
    const int64_t end = (beg + chunk < n) ? beg + chunk : n;
    double acc = 0.0;
    int zero = 0;
    LogProduct lp;

    // Coalesced 16-byte loads: two RTs per thread per step.  A slice that starts on an odd element (caller passed
    // an 8-byte-aligned view) peels one head element and an odd remainder leaves one tail element; both are handed
    // to single threads in a final pass through the same (single) evaluation site.
    int64_t b = beg;
    const bool head = (reinterpret_cast<uintptr_t>(x + b) & 15) && b < end;
    if (head) b++;
    const int64_t npair = (end - b) >> 1;
    const bool tail = (end > b) && ((end - b) & 1);
    const double2* x2 = reinterpret_cast<const double2*>(x + b);
    // CTA-uniform trip count (every thread runs the same number of rounds, the last ones possibly empty-handed), so
    // that the parking lot can be drained by the whole CTA between rounds before it overflows.
    // one_each (small grids, where latency counts rather than throughput): one RT per thread and round, 8-byte loads
    // In that mode the slices of a set interleave (slice s takes trials s, s + slices, ...): trials that need refinement
    // are neighbours in the |rt|-sorted array, and this spreads them over all the CTAs of the set.
    // (32-bit counters: a slice never holds 2^30 RTs, see plan_slices)
    const int mine = (n > s) ? (int)((n - s + slices - 1) / slices) : 0;
    const int npair32 = (int)npair;
    const int rounds = one_each ? (mine + kBlock - 1) / kBlock
                                : (npair32 + kBlock - 1) / kBlock + 1;   // + 1: the head / tail elements
    double acc_slow = 0.0;
    auto drain = [&](int parked) {
        if (Impl::coop_applies(set, parked)) {   // uniform over the CTA
            // a handful of parked trials: one warp each, the warp's lanes sharing the trial's quadrature nodes
#pragma unroll 1
            for (int q = threadIdx.x >> 5; q < parked; q += kBlock / 32) {
                const double p0 = mix_outlier(Impl::density_coop(Impl::parked(set, q), set), omp, wp);
                if ((threadIdx.x & 31) == 0) {
                    zero |= (p0 == 0);
                    acc_slow += log(p0);
                }
            }
            return;
        }
#pragma unroll 1
        for (int q = threadIdx.x; q < parked; q += kBlock) {
            const double p0 = mix_outlier(Impl::density_generic(Impl::parked(set, q), set, ctx), omp, wp);
            zero |= (p0 == 0);
            acc_slow += log(p0);
        }
    };
    if (Impl::whole_set_generic(set)) {
        // No table covers this parameter set: every trial goes through the operation-for-operation evaluator.  Kept apart
        // from the loop below so that the hot loop of the fixed-grid instantiation contains no call at all (a call site
        // there makes the register allocation of the whole loop depend on the callee).
#pragma unroll 1
        for (int64_t i = beg + threadIdx.x; i < end; i += kBlock) {
            const double p0 = mix_outlier(Impl::density_generic(x[i], set, ctx), omp, wp);
            zero |= (p0 == 0);
            acc_slow += log(p0);
        }
    } else {
#pragma unroll 1
    for (int r = 0; r < rounds; r++) {
        double v0 = 0.0, v1 = 0.0;
        int cnt = 0;
        const int i = r * kBlock + threadIdx.x;
        if (one_each) {
            if (i < mine) { v0 = x[(int64_t)i * slices + s]; cnt = 1; }
        } else if (i < npair32) {
            const double2 xx = __ldg(x2 + i);
            v0 = xx.x;
            v1 = xx.y;
            cnt = 2;
        } else if (r == rounds - 1) {
            if (head && threadIdx.x == 0) { v0 = x[beg]; cnt = 1; }
            if (tail && threadIdx.x == kBlock - 1) { v0 = x[end - 1]; cnt = 1; }
        }
#pragma unroll 1
        for (int h = 0; h < cnt; h++) {
            const double xv = h ? v1 : v0;
            bool defer;
            double dens = Impl::template density<ADAPT>(xv, set, ctx, &defer);
            if (ADAPT && defer) {
                // park the trial; evaluated by the whole CTA between rounds.  The lot is drained whenever fewer than one
                // round's worth of slots (two per thread) is left, so it cannot be full here; should that invariant ever
                // be broken the kernel stops loudly instead of evaluating inline (a call site in this loop costs the
                // common path registers) or dropping the trial.
                if (!Impl::park(set, xv)) __trap();
                continue;
            }
            lp.add(mix_outlier(dens, omp, wp), zero);
        }
        if (ADAPT && Impl::kParks && Impl::may_park(set) && (r % kParkCheck) == kParkCheck - 1) {   // uniform over the CTA
            // Room for kParkCheck more rounds (two trials per thread each)?  If not, everybody evaluates the parked trials
            // now.  Looking only every kParkCheck-th round keeps the barrier this needs out of most rounds.
            const int parked = Impl::parked_count(set);   // barrier inside
            if (parked > Impl::kParkCapacity - kParkCheck * 2 * kBlock) {
                drain(parked);
                Impl::parked_reset(set);                   // barriers inside
            }
        }
    }
    }
    // remaining parked trials: dense pass through the generic evaluator
    if (ADAPT && Impl::kParks && Impl::may_park(set)) drain(Impl::parked_count(set));
    acc = lp.value() + acc_slow;

    int any_zero;
    const double tot = block_sum_or<kBlock>(acc, zero, &any_zero);
    if (threadIdx.x == 0) {
        if (out_zero) {
            out_sum[bid] = tot;
            out_zero[bid] = any_zero;
            if (done) {
                // fused finalize: whoever finishes last adds the set's partials up in slice order (same order and
                // result as finalize_kernel, whatever the arrival order) and leaves the counter at zero for the next launch
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
            // wfpt.pyx:71-72: the first p == 0 returns -inf (a NaN elsewhere never gets to overrule it)
            out_sum[bid] = any_zero ? -CUDART_INF : tot;
        }
    }
}

// Sequential (fixed-order) sum of the per-slice partials: deterministic and independent of launch geometry.
__global__ void finalize_kernel(const double* __restrict__ part_sum, const int* __restrict__ part_zero, int64_t P,
                                int slices, int early_exit, double* __restrict__ out);

template <class Impl, bool ADAPT>
__global__ void __launch_bounds__(kBlock) pdf_array_kernel(const double* __restrict__ x, int64_t n, Params prm, Opts o,
                                                         double w_outlier, int logp, double* __restrict__ y,
                                                         const __grid_constant__ typename Impl::Ctx ctx)
{
    const int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x;
    typename Impl::Set set;
    Impl::prepare(set, prm, o, ctx);
    if (i >= n) return;
    // wfpt.pyx:44: y * (1 - p_outlier) + (w_outlier * p_outlier), then optionally np.log
    const double wp = __dmul_rn(w_outlier, prm.p_outlier);
    const double omp = __dsub_rn(1.0, prm.p_outlier);
    bool defer = true;
    double dens = 0.0;
    if (!Impl::whole_set_generic(set)) dens = Impl::template density<ADAPT>(x[i], set, ctx, &defer);
    if (defer) dens = Impl::density_generic(x[i], set, ctx);
    const double p = mix_outlier(dens, omp, wp);
    y[i] = logp ? log(p) : p;
}

template <class Impl>
__global__ void __launch_bounds__(kBlock) like_multi_kernel(const double* __restrict__ x, int64_t n, MultiSrc src,
                                                          Opts o, double p_outlier, double w_outlier,
                                                          const double* __restrict__ vscale,
                                                          const uint8_t* __restrict__ skip, int code999,
                                                          double* __restrict__ out_sum, int* __restrict__ out_zero,
                                                          const __grid_constant__ typename Impl::Ctx ctx)
{
    const double wp = __dmul_rn(w_outlier, p_outlier);
    const double omp = __dsub_rn(1.0, p_outlier);
    Impl::prepare_multi(ctx);   // CTA-collective: shared tables of the per-trial evaluator (fast variant)
    double acc = 0.0;
    int zero = 0;
    const int64_t stride = (int64_t)gridDim.x * kBlock;
    for (int64_t i = (int64_t)blockIdx.x * kBlock + threadIdx.x; i < n; i += stride) {
        if (skip && skip[i]) continue;
        double q[7];
#pragma unroll
        for (int k = 0; k < 7; k++) q[k] = src.ptr[k] ? __ldg(src.ptr[k] + i) : src.scalar[k];
        if (vscale) q[0] = __dmul_rn(vscale[i], q[0]);  // wfpt.pyx:133: (qs[1]-qs[0]) * v
        const double xi = x[i];
        double p;
        if (!code999 || fabs(xi) != 999.) {
            Params prm{q[0], q[1], q[2], q[3], q[4], q[5], q[6], p_outlier};
            p = mix_outlier(Impl::density_once(xi, prm, o, ctx), omp, wp);
        } else if (xi == 999.) {
            p = Impl::prob_ub(q[0], q[2], q[3]);
        } else {
            p = 1 - Impl::prob_ub(q[0], q[2], q[3]);
        }
        zero |= (p == 0);
        acc += log(p);
    }
    int any_zero;
    const double tot = block_sum_or<kBlock>(acc, zero, &any_zero);
    if (threadIdx.x == 0) {
        out_sum[blockIdx.x] = tot;
        out_zero[blockIdx.x] = any_zero;
    }
}

}  // namespace wfpt
