// Shared declarations for the sm_100a WFPT / simulator library (internal; the public C ABI is include/wfpt_b200.h).
#pragma once
#include <cuda_runtime.h>
#include <math_constants.h>
#include <stdint.h>
#include <stddef.h>

namespace wfpt {

// Integration / series options of one call; mirrors the keyword block every reference entry point carries
// (/root/reference/src/wfpt.pyx:32-34,54-56; /root/reference/hddm/likelihoods.py:105-113).
struct Opts {
    double err;        // series truncation bound
    double simps_err;  // adaptive Simpson tolerance
    int n_st;          // fixed grid: intervals over t; adaptive: max recursion depth over t
    int n_sz;          // same for z
    int use_adaptive;
};

// One parameter set of the batched likelihood: row layout of params[P][8].
struct Params {
    double v, sv, a, z, sz, t, st, p_outlier;
};

// Per-trial parameter sources for wiener_like_multi (NULL pointer -> use the scalar).
struct MultiSrc {
    const double* ptr[7];  // v, sv, a, z, sz, t, st
    double scalar[7];
};

// One parameter row carried in the kernel arguments (P == 1 calls from the host).
struct InlineRow {
    double v[8];
    int use;
};

constexpr int kBlock = 128;          // threads per CTA for the likelihood kernels
constexpr int kMaxDepth = 24;        // deepest adaptive-Simpson recursion accepted (2^24 nodes is already absurd)
constexpr double kPi = 3.14159265358979323846;       // M_PI
constexpr double kPiSq = 9.869604401089358;          // fl(M_PI*M_PI) == pow(M_PI, 2.0)

// ---- launchers implemented in like_faithful.cu / like_fast.cu -----------------------------------------------------
struct BatchArgs {
    const double* x;      // [n] signed RTs (device)
    int64_t n;
    const double* params; // [P][8] (device)
    int64_t P;
    Opts o;
    double w_outlier;
    int slices;           // trial slices per parameter set (>=1)
    int64_t chunk;        // trials per slice
    double* part_sum;     // [P*slices] when slices>1, else == out
    int* part_zero;       // [P*slices] when slices>1, else nullptr
    double* out;          // [P]
    cudaStream_t stream;
    const int64_t* seg = nullptr;  // optional [P][2] (device): set p is evaluated on x[seg[p][0] .. seg[p][1]) instead of all of x
    int one_each = 0;              // 1: chunk may be as small as one RT per thread (small grids; see plan_slices)
    double* park_x = nullptr;      // adaptive launches: [P*slices][park_cap] trials parked for the refinement kernel, and
    int* park_n = nullptr;         // [P*slices] their counts (library scratch; without them the one-kernel form runs)
    int park_cap = 0;
    const double* host_row = nullptr;  // P == 1 and the caller holds the row on the host: passed in the kernel arguments
    int* done = nullptr;           // optional [P] zeroed counters (device): when given (slices > 1), the slice that finishes last
                                   // sums its set's partials in slice order into out[p] and re-zeroes the counter, so that no
                                   // separate finalize launch is needed
};

inline InlineRow inline_row(const BatchArgs& a)
{
    InlineRow r;
    r.use = (a.host_row != nullptr && a.P == 1) ? 1 : 0;
    for (int k = 0; k < 8; k++) r.v[k] = r.use ? a.host_row[k] : 0.0;
    return r;
}

cudaError_t launch_like_batch_faithful(const BatchArgs& a);
cudaError_t launch_like_batch_fast(const BatchArgs& a);
struct KTable;
bool lean_applies(const BatchArgs& a, const KTable& kt);           // like_lean.cu: fixed-grid launches with n_sz = 2
bool lean_adapt_applies(const BatchArgs& a, const KTable& kt);     // like_lean.cu: adaptive launches on one RT array
cudaError_t launch_like_lean(const BatchArgs& a, const KTable& kt);

cudaError_t launch_kflip_census(const double* x, int64_t n, const double* params, int64_t P, const Opts& o,
                                unsigned long long* counts, cudaStream_t stream);

struct PdfArrayArgs {
    const double* x; int64_t n; Params prm; Opts o; double w_outlier; int logp; double* y; cudaStream_t stream;
};
cudaError_t launch_pdf_array_faithful(const PdfArrayArgs& a);
cudaError_t launch_pdf_array_fast(const PdfArrayArgs& a);

struct MultiArgs {
    const double* x; int64_t n; MultiSrc src; Opts o; double p_outlier; double w_outlier;
    const double* vscale;   // optional [n]: v_i is multiplied by vscale[i] (RL-DDM drift scaling), or nullptr
    const uint8_t* skip;    // optional [n]: 1 -> trial contributes nothing (first trial of an RL group)
    int early_exit;         // 1: any p==0 -> -inf (wiener_like_rlddm); 0: plain IEEE sum (wiener_like_multi)
    int code999;            // 1: |x|==999 means "choice only" (wiener_like_multi)
    int nblocks;            // CTAs; partials has nblocks entries
    double* part_sum; int* part_zero; double* out; cudaStream_t stream;
};
cudaError_t launch_like_multi_faithful(const MultiArgs& a);
cudaError_t launch_like_multi_fast(const MultiArgs& a);

// ---- hooks exported by api.cu for sibling translation units -------------------------------------------------------------
int api_fail(const char* msg);
int api_sim_scratch(int which, size_t bytes, void** out);   // grow-only device buffers of the simulator entry points (0..5)
int api_sm_count();
int api_scratch_pair(size_t bytes0, size_t bytes1, void** p0, void** p1);   // two grow-only staging buffers (kStageX / kStageOut)
// Held for the whole body of every `_host` entry point (stage in, launch, synchronise, copy out): those calls share one
// pinned staging block, the grow-only scratch and stream 0 per device, so two host threads inside them at once would
// overwrite each other's inputs and read each other's results.  Per device, re-entrant (host entry points call each other).
struct HostCall {
    HostCall();
    ~HostCall();
    HostCall(const HostCall&) = delete;
    HostCall& operator=(const HostCall&) = delete;
private:
    int dev_;
};
int api_check_opts(double err, int n_st, int n_sz, int use_adaptive, double simps_err, int variant, Opts* o);
void api_plan_slices(int64_t n, int64_t P, int* slices, int64_t* chunk);
cudaError_t api_launch_finalize(const double* part_sum, const int* part_zero, int64_t P, int slices, int early_exit,
                                double* out, cudaStream_t stream);
void count_launch(int k);

// ---- warp/CTA reductions ------------------------------------------------------------------------------------------------
__device__ __forceinline__ double warp_sum(double v)
{
#pragma unroll
    for (int off = 16; off > 0; off >>= 1) v += __shfl_down_sync(0xffffffffu, v, off);
    return v;
}

// Sum over the CTA (result valid in thread 0) and OR of `flag` over the CTA (valid in all threads).
template <int BLOCK>
__device__ __forceinline__ double block_sum_or(double v, int flag, int* any_flag)
{
    __shared__ double s_part[BLOCK / 32];
    const int lane = threadIdx.x & 31, wid = threadIdx.x >> 5;
    v = warp_sum(v);
    if (lane == 0) s_part[wid] = v;
    *any_flag = __syncthreads_or(flag);
    double tot = 0.0;
    if (threadIdx.x == 0) {
#pragma unroll
        for (int i = 0; i < BLOCK / 32; i++) tot += s_part[i];
    }
    return tot;
}

}  // namespace wfpt
