// wfpt_like_batch_noalloc: the batched likelihood for callers that must not allocate device memory.
//
// Who that is: a sampler or service thread that runs inside a process whose GPU memory already belongs to someone
// else (a framework's caching pool that has reserved the card, a neighbour job on a shared GPU), where a fresh
// cudaMalloc can fail long after start-up; a caller on a latency budget, for whom cudaMalloc's implicit
// device-wide synchronisation stalls every other stream; and code that issues work from a context in which the
// runtime forbids allocation calls altogether.  For them the per-slice partial sums of the two-stage trial
// reduction, which wfpt_like_batch keeps in library-managed (cudaMalloc'ed, grow-only) scratch, live here in a
// fixed array inside the library image instead.
//
// Cost and limits: every GPU that loads libwfpt_b200.so carries the array (kNoAllocSlots * 12 bytes = 192 KiB) until
// the process exits.  A call needs P * slices slots, where `slices` comes from the library's general slicing rule
// (api_plan_slices) WITHOUT the small-grid re-plan wfpt_like_batch adds for latency-bound grids: for those the two
// entry points cut the trials differently and agree to rounding, not to the last bit.  Likewise large launches, for which
// wfpt_like_batch may pick kernels that keep lists in further (allocated) scratch (adaptive integration): this entry point
// stays with kernels whose only scratch is the array below.  Calls needing more than
// kNoAllocSlots are refused with an error, never truncated.
//
// Sharing: all calls on a device use the same array and it still holds the previous call's partials on entry.
// Nothing in it has to start at zero -- the first-stage kernel overwrites every slot the second stage reads -- so no
// clearing is done.  Two calls whose kernels overlap on the GPU (different streams, or different host threads) race
// on the array and corrupt each other's sums: serialise them.
//
// Set-up: the device address of the array is looked up (cudaGetSymbolAddress, with that device current) by the first
// call on each device and cached.  Make that first call on each device before any stream capture begins.  The cached
// address belongs to the device's primary context: after cudaDeviceReset it is stale, and this entry point must not be
// used again in that process.
#include <cstdio>
#include <mutex>

#include "../../include/wfpt_b200.h"
#include "common.cuh"

namespace wfpt {

constexpr int64_t kNoAllocSlots = WFPT_NOALLOC_MAX_SLOTS;
constexpr int kNoAllocMaxDev = 16;

// The scratch itself.  In host code these names are only handles for cudaGetSymbolAddress, never device pointers.
static __device__ __align__(256) double g_noalloc_sum[kNoAllocSlots];
static __device__ __align__(256) int g_noalloc_zero[kNoAllocSlots];

static std::mutex g_noalloc_mu;
static double* g_sum_ptr[kNoAllocMaxDev] = {nullptr};
static int* g_zero_ptr[kNoAllocMaxDev] = {nullptr};

// Device addresses of the scratch on the current device (looked up once per device).
static int noalloc_scratch(double** sum, int** zero)
{
    int dev = 0;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return api_fail(cudaGetErrorString(e));
    if (dev < 0 || dev >= kNoAllocMaxDev) return api_fail("wfpt_like_batch_noalloc: device index beyond the 16 this library tracks");
    std::lock_guard<std::mutex> lk(g_noalloc_mu);
    if (!g_sum_ptr[dev] || !g_zero_ptr[dev]) {
        void *ps = nullptr, *pz = nullptr;
        e = cudaGetSymbolAddress(&ps, g_noalloc_sum);
        if (e != cudaSuccess) return api_fail(cudaGetErrorString(e));
        e = cudaGetSymbolAddress(&pz, g_noalloc_zero);
        if (e != cudaSuccess) return api_fail(cudaGetErrorString(e));
        g_sum_ptr[dev] = static_cast<double*>(ps);
        g_zero_ptr[dev] = static_cast<int*>(pz);
    }
    *sum = g_sum_ptr[dev];
    *zero = g_zero_ptr[dev];
    return 0;
}

}  // namespace wfpt

using namespace wfpt;

extern "C" int wfpt_like_batch_noalloc(const double* d_x, int64_t n, const double* d_params, int64_t P, double err,
                                       int n_st, int n_sz, int use_adaptive, double simps_err, double w_outlier,
                                       int variant, double* d_out, void* stream)
{
    Opts o;
    if (api_check_opts(err, n_st, n_sz, use_adaptive, simps_err, variant, &o)) return 1;
    if (n < 0 || P < 0) return api_fail("wfpt_like_batch_noalloc: negative size");
    if (P == 0) return 0;
    if (!d_params || !d_out || (n > 0 && !d_x)) return api_fail("wfpt_like_batch_noalloc: null pointer argument");
    if ((reinterpret_cast<uintptr_t>(d_x) & 7) != 0) return api_fail("wfpt_like_batch_noalloc: d_x must be 8-byte aligned");

    BatchArgs a;
    a.x = d_x; a.n = n; a.params = d_params; a.P = P; a.o = o; a.w_outlier = w_outlier; a.out = d_out;
    a.stream = (cudaStream_t)stream;
    api_plan_slices(n, P, &a.slices, &a.chunk);
    const double need = (double)P * (double)a.slices;
    if (need > (double)kNoAllocSlots) {
        char msg[256];
        snprintf(msg, sizeof(msg),
                 "wfpt_like_batch_noalloc: %lld parameter sets x %d trial slices need %.0f scratch slots, the built-in "
                 "scratch holds %lld; split the batch or use wfpt_like_batch",
                 (long long)P, a.slices, need, (long long)kNoAllocSlots);
        return api_fail(msg);
    }
    if (noalloc_scratch(&a.part_sum, &a.part_zero)) return 1;

    // Two steps for every geometry: the first-stage kernel fills the scratch, the second stage reads it in a fixed
    // order and writes d_out.  With several slices per set the scratch holds (partial sum, saw-a-zero flag) pairs;
    // with a single slice the kernel's own epilogue has already applied the "-inf on the first zero density" rule,
    // so the finished per-set value is routed through the scratch and the flags are not consulted.
    BatchArgs k = a;
    if (a.slices == 1) k.out = a.part_sum;
    cudaError_t e = (variant == WFPT_VARIANT_FAITHFUL) ? launch_like_batch_faithful(k) : launch_like_batch_fast(k);
    if (e != cudaSuccess) return api_fail(cudaGetErrorString(e));
    e = api_launch_finalize(a.part_sum, a.part_zero, P, a.slices, a.slices > 1 ? 1 : 0, d_out, a.stream);
    if (e != cudaSuccess) return api_fail(cudaGetErrorString(e));
    count_launch(2);
    return 0;
}
