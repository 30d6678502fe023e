// Fast-variant instantiations of the likelihood kernels (table-driven evaluator, wfpt_fast.cuh).
#include <cstdlib>

#include "like_kernels.cuh"
#include "wfpt_fast.cuh"

namespace wfpt {

struct FastImpl {
    using Ctx = KTable;
    using Set = fast::Set;
    __device__ static __forceinline__ void prepare(Set& s, const Params& prm, const Opts& o, const Ctx& kt)
    {
        fast::prepare(s, prm, o, kt);
    }
    template <bool ADAPT>
    __device__ static __forceinline__ double density(double x, Set& s, const Ctx&, bool* defer)
    {
        return fast::density<ADAPT>(x, s, defer);
    }
    __device__ static __forceinline__ double density_generic(double x, const Set& s, const Ctx&)
    {
        return fast::density_generic(x, s);
    }
    // parameter sets without a table (err / layout not covered): the kernels route every trial to density_generic
    __device__ static __forceinline__ bool whole_set_generic(const Set& s) { return s.mode == fast::kModeGeneric; }
    // Few parked trials: one warp per trial instead of one thread per trial (see fast::refine_coop).  Every thread of
    // the CTA calls it; returns the density of parked trial q to every lane of the warp that took it.
    __device__ static __forceinline__ bool coop_applies(const Set& s, int parked)
    {
        return parked <= 32 && s.mode == fast::kModeAdaptive && s.kind != 4 && s.tab->depth_t == 2;
    }
    __device__ static __forceinline__ double density_coop(double x, const Set& s)
    {
        return fast::refine_coop(x, s.tab, s.kind, s.kl_max);
    }
    static constexpr bool kParks = true;
    static constexpr int kParkCapacity = fast::kQueue;
    static_assert(fast::kQueue > kParkCheck * 2 * kBlock, "parking lot smaller than what can arrive between two checks");
    // only adaptive (refinement) and table-less parameter sets ever defer a trial
    __device__ static __forceinline__ bool may_park(const Set& s) { return s.mode == fast::kModeAdaptive; }
    __device__ static __forceinline__ void parked_reset(const Set&)
    {
        __syncthreads();
        if (threadIdx.x == 0) fast::tables().qcount = 0;
        __syncthreads();
    }
    // Parking lot in shared memory (fast::Tables): returns false when full.
    __device__ static __forceinline__ bool park(Set& s, double x)
    {
        fast::Tables& T = fast::tables();
        const int slot = atomicAdd(&T.qcount, 1);
        if (slot >= fast::kQueue) return false;
        fast::parking_lot()[slot] = x;
        return true;
    }
    __device__ static __forceinline__ int parked_count(const Set&)
    {
        __syncthreads();
        const int n = fast::tables().qcount;
        return n < fast::kQueue ? n : fast::kQueue;
    }
    __device__ static __forceinline__ double parked(const Set&, int q) { return fast::parking_lot()[q]; }
    __device__ static __forceinline__ void prepare_multi(const Ctx& kt) { fast::prepare_series_only(kt); }
    __device__ static __forceinline__ double density_once(double x, const Params& prm, const Opts& o, const Ctx& kt)
    {
        return fast::density_once(x, prm, o, kt);
    }
    __device__ static __forceinline__ double prob_ub(double v, double a, double z) { return faithful::prob_ub(v, a, z); }
};

cudaError_t launch_like_batch_fast(const BatchArgs& a)
{
    const int64_t nblk = a.P * (int64_t)a.slices;
    if (nblk == 0) return cudaSuccess;
    {
        const KTable kt = get_ktable(a.o.err);
        if ((lean_applies(a, kt) || lean_adapt_applies(a, kt)) && !getenv("WFPT_NO_LEAN")) return launch_like_lean(a, kt);
    }
    if (a.o.use_adaptive)
        like_batch_kernel<FastImpl, true><<<(unsigned)nblk, kBlock, 0, a.stream>>>(
            a.x, a.n, a.params, a.P, a.o, a.w_outlier, a.slices, a.chunk, a.slices > 1 ? a.part_sum : a.out,
            a.slices > 1 ? a.part_zero : nullptr, a.seg, a.slices > 1 ? a.done : nullptr, a.out, a.one_each, inline_row(a), get_ktable(a.o.err));
    else
        like_batch_kernel<FastImpl, false><<<(unsigned)nblk, kBlock, 0, a.stream>>>(
            a.x, a.n, a.params, a.P, a.o, a.w_outlier, a.slices, a.chunk, a.slices > 1 ? a.part_sum : a.out,
            a.slices > 1 ? a.part_zero : nullptr, a.seg, a.slices > 1 ? a.done : nullptr, a.out, a.one_each, inline_row(a), get_ktable(a.o.err));
    return cudaGetLastError();
}

// Census of series-term decisions (SURVEY 7 "measure, don't assume"): the fast evaluator decides the expansion and K on
// tt = x'/a^2 formed from a reciprocal and one residual correction (fast::make_tnode), the reference on the IEEE quotient
// x' / pow(a, 2) (/root/reference/src/pdf.pxi:35-60).  For every (parameter set, trial, staged t-node) this kernel forms
// both, looks both up in the same table and counts (0) the pairs visited, (1) pairs whose two tt differ in any bit,
// (2) pairs whose descriptors differ -- a "K flip", which would move that density by up to `err`.  One CTA per set.
__global__ void __launch_bounds__(kBlock) kflip_census_kernel(const double* __restrict__ x, int64_t n,
                                                             const double* __restrict__ params, int64_t P, Opts o,
                                                             unsigned long long* __restrict__ counts,
                                                             const __grid_constant__ KTable kt)
{
    const int64_t p = blockIdx.x;
    if (p >= P) return;
    const double* row = params + 8 * p;
    Params prm{row[0], row[1], row[2], row[3], row[4], row[5], row[6], row[7]};
    fast::Set set;
    fast::prepare(set, prm, o, kt);
    if (set.mode != fast::kModePoint && set.mode != fast::kModeFixed && set.mode != fast::kModeAdaptive) return;
    const fast::Tables& T = *set.tab;
    const double a2_ref = __dmul_rn(prm.a, prm.a);   // pow(a, 2.0) is a*a correctly rounded (checked on the host for every a)
    unsigned long long seen = 0, tt_diff = 0, flips = 0;
    for (int64_t i = threadIdx.x; i < n; i += kBlock) {
        const double ax = fabs(x[i]);
        if ((ax - set.tlo) < 0) continue;
        for (int k = 0; k <= set.nt; k++) {
            const double xp = ax - T.tn[k];
            if (!(xp > 0)) continue;
            const double q0 = xp * set.inv_a2;
            const double tt_fast = fma(fma(-q0, set.a2, xp), set.inv_a2, q0);
            const double tt_ref = __ddiv_rn(xp, a2_ref);
            seen++;
            tt_diff += (__double_as_longlong(tt_fast) != __double_as_longlong(tt_ref));
            flips += (fast::klookup_scan(kt, tt_fast) != fast::klookup_scan(kt, tt_ref));
        }
    }
    for (int off = 16; off > 0; off >>= 1) {
        seen += __shfl_down_sync(0xffffffffu, seen, off);
        tt_diff += __shfl_down_sync(0xffffffffu, tt_diff, off);
        flips += __shfl_down_sync(0xffffffffu, flips, off);
    }
    if ((threadIdx.x & 31) == 0) {
        atomicAdd(&counts[0], seen);
        atomicAdd(&counts[1], tt_diff);
        atomicAdd(&counts[2], flips);
    }
}

cudaError_t launch_kflip_census(const double* x, int64_t n, const double* params, int64_t P, const Opts& o,
                                unsigned long long* counts, cudaStream_t stream)
{
    if (P == 0) return cudaSuccess;
    kflip_census_kernel<<<(unsigned)P, kBlock, 0, stream>>>(x, n, params, P, o, counts, get_ktable(o.err));
    return cudaGetLastError();
}

cudaError_t launch_pdf_array_fast(const PdfArrayArgs& a)
{
    if (a.n == 0) return cudaSuccess;
    const unsigned nblk = (unsigned)((a.n + kBlock - 1) / kBlock);
    if (a.o.use_adaptive)
        pdf_array_kernel<FastImpl, true><<<nblk, kBlock, 0, a.stream>>>(a.x, a.n, a.prm, a.o, a.w_outlier, a.logp, a.y,
                                                                       get_ktable(a.o.err));
    else
        pdf_array_kernel<FastImpl, false><<<nblk, kBlock, 0, a.stream>>>(a.x, a.n, a.prm, a.o, a.w_outlier, a.logp, a.y,
                                                                        get_ktable(a.o.err));
    return cudaGetLastError();
}

cudaError_t launch_like_multi_fast(const MultiArgs& a)
{
    like_multi_kernel<FastImpl><<<(unsigned)a.nblocks, kBlock, 0, a.stream>>>(
        a.x, a.n, a.src, a.o, a.p_outlier, a.w_outlier, a.vscale, a.skip, a.code999, a.part_sum, a.part_zero,
        get_ktable(a.o.err));
    return cudaGetLastError();
}

}  // namespace wfpt
