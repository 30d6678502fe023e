// Faithful-variant instantiations of the likelihood kernels.  THIS FILE IS COMPILED WITH --fmad=false (see
// hddm_b200/build.py): see wfpt_faithful.cuh for why.
#include "like_kernels.cuh"
#include "wfpt_faithful.cuh"

namespace wfpt {

struct FaithfulImpl {
    struct Ctx { int unused; };
    struct Set {
        Params prm;
        Opts o;
    };
    __device__ static __forceinline__ void prepare(Set& s, const Params& prm, const Opts& o, const Ctx&)
    {
        s.prm = prm;
        s.o = o;
    }
    template <bool ADAPT>
    __device__ static __forceinline__ double density(double x, const Set& s, const Ctx&, bool* defer)
    {
        *defer = false;
        return faithful::full_pdf(x, s.prm.v, s.prm.sv, s.prm.a, s.prm.z, s.prm.sz, s.prm.t, s.prm.st, s.o);
    }
    __device__ static __forceinline__ double density_generic(double x, const Set& s, const Ctx& c)
    {
        bool d;
        return density<true>(x, s, c, &d);
    }
    __device__ static __forceinline__ bool whole_set_generic(const Set&) { return false; }
    __device__ static __forceinline__ bool coop_applies(const Set&, int) { return false; }
    __device__ static __forceinline__ double density_coop(double, const Set&) { return 0.0; }
    static constexpr bool kParks = false;
    static constexpr int kParkCapacity = 0;
    __device__ static __forceinline__ bool may_park(const Set&) { return false; }
    __device__ static __forceinline__ void parked_reset(const Set&) {}
    __device__ static __forceinline__ bool park(Set&, double) { return false; }
    __device__ static __forceinline__ int parked_count(const Set&) { return 0; }
    __device__ static __forceinline__ double parked(const Set&, int) { return 0.0; }
    __device__ static __forceinline__ void prepare_multi(const Ctx&) {}
    __device__ static __forceinline__ double density_once(double x, const Params& prm, const Opts& o, const Ctx&)
    {
        return faithful::full_pdf(x, prm.v, prm.sv, prm.a, prm.z, prm.sz, prm.t, prm.st, o);
    }
    __device__ static __forceinline__ double prob_ub(double v, double a, double z) { return faithful::prob_ub(v, a, z); }
};

cudaError_t launch_like_batch_faithful(const BatchArgs& a)
{
    const int64_t nblk = a.P * (int64_t)a.slices;
    if (nblk == 0) return cudaSuccess;
    like_batch_kernel<FaithfulImpl, false><<<(unsigned)nblk, kBlock, 0, a.stream>>>(
        a.x, a.n, a.params, a.P, a.o, a.w_outlier, a.slices, a.chunk, a.slices > 1 ? a.part_sum : a.out,
        a.slices > 1 ? a.part_zero : nullptr, a.seg, a.slices > 1 ? a.done : nullptr, a.out, a.one_each, inline_row(a), FaithfulImpl::Ctx{0});
    return cudaGetLastError();
}

cudaError_t launch_pdf_array_faithful(const PdfArrayArgs& a)
{
    if (a.n == 0) return cudaSuccess;
    const unsigned nblk = (unsigned)((a.n + kBlock - 1) / kBlock);
    pdf_array_kernel<FaithfulImpl, false><<<nblk, kBlock, 0, a.stream>>>(a.x, a.n, a.prm, a.o, a.w_outlier, a.logp, a.y, FaithfulImpl::Ctx{0});
    return cudaGetLastError();
}

cudaError_t launch_like_multi_faithful(const MultiArgs& a)
{
    like_multi_kernel<FaithfulImpl><<<(unsigned)a.nblocks, kBlock, 0, a.stream>>>(
        a.x, a.n, a.src, a.o, a.p_outlier, a.w_outlier, a.vscale, a.skip, a.code999, a.part_sum, a.part_zero, FaithfulImpl::Ctx{0});
    return cudaGetLastError();
}

}  // namespace wfpt
