// Fast-variant instantiations of the likelihood kernels (table-driven evaluator, wfpt_fast.cuh).
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
        T.queue[slot] = x;
        return true;
    }
    __device__ static __forceinline__ int parked_count(const Set&)
    {
        __syncthreads();
        const int n = fast::tables().qcount;
        return n < fast::kQueue ? n : fast::kQueue;
    }
    __device__ static __forceinline__ double parked(const Set&, int q) { return fast::tables().queue[q]; }
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
    if (a.o.use_adaptive)
        like_batch_kernel<FastImpl, true><<<(unsigned)nblk, kBlock, 0, a.stream>>>(
            a.x, a.n, a.params, a.P, a.o, a.w_outlier, a.slices, a.chunk, a.slices > 1 ? a.part_sum : a.out,
            a.slices > 1 ? a.part_zero : nullptr, a.seg, a.slices > 1 ? a.done : nullptr, a.out, a.one_each, get_ktable(a.o.err));
    else
        like_batch_kernel<FastImpl, false><<<(unsigned)nblk, kBlock, 0, a.stream>>>(
            a.x, a.n, a.params, a.P, a.o, a.w_outlier, a.slices, a.chunk, a.slices > 1 ? a.part_sum : a.out,
            a.slices > 1 ? a.part_zero : nullptr, a.seg, a.slices > 1 ? a.done : nullptr, a.out, a.one_each, get_ktable(a.o.err));
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
