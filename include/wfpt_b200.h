/*
 * wfpt_b200.h -- C ABI of libwfpt_b200.so: the WFPT log-likelihood and the Euler-Maruyama sequential-sampling
 * simulators of HDDM as hand-written sm_100a CUDA kernels.
 *
 * The reference has no FFI layer for this path: its "operator API" is the set of module-level callables of the
 * Cython extension `wfpt` and of `hddm.simulators.simulator`.  Each entry point below names the reference
 * callable it computes (paths relative to /root/reference).  The Python drop-in (hddm_b200/wfpt.py, top-level
 * wfpt.py, hddm_b200/simulators.py) binds these with ctypes; INTEGRATION.md shows the stub.
 *
 * Conventions
 *   - every function returns 0 on success, non-zero on failure; wfpt_last_error() then describes the failure
 *     (thread-local string).  Out-of-support MODEL parameters are not failures: they produce -inf / NaN /
 *     N*log(w_outlier*p_outlier) exactly as the reference does.
 *   - `d_` pointers are device pointers on the current CUDA device; `h_` pointers are host pointers.  Functions
 *     without a `_host` suffix take device pointers, enqueue on `stream` (a cudaStream_t passed as void*, NULL =
 *     default stream) and return without synchronising.  `_host` functions copy in, run, copy out and synchronise.
 *   - all likelihood arithmetic is IEEE binary64.  RTs are signed: positive = upper boundary, negative = lower
 *     (hddm/utils.py:489-517 flip_errors).
 *   - parameter-set rows are double[8] = {v, sv, a, z, sz, t, st, p_outlier}.
 *   - integration options, as in every reference signature: err, n_st, n_sz, use_adaptive, simps_err
 *     (src/pdf.pxi:104-108).  With use_adaptive != 0, n_st/n_sz are recursion depths and must be <= 24.
 *   - threads and streams: `_host` functions hold a per-device lock for their whole body (they share one staging block,
 *     the library's scratch and stream 0), so they may be called from several host threads; calls on one device simply
 *     run one after another.  Device-pointer functions are asynchronous and share, per device, the scratch that holds the
 *     per-slice partial sums and the arrival counters of the trial reduction (wfpt_like_batch / wfpt_like_segments whenever a
 *     set's trials are cut into more than one slice, wfpt_like_multi always): two such calls whose kernels can overlap on the
 *     GPU (different streams, or different host threads) race on it -- enqueue them on one stream or order them with events.
 *   - `variant`: WFPT_VARIANT_FAST (default product path) or WFPT_VARIANT_FAITHFUL (operation-for-operation
 *     evaluation without FMA contraction; slower; kept as the on-device cross-check).
 */
#ifndef WFPT_B200_H
#define WFPT_B200_H

#include <stddef.h>
#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define WFPT_VARIANT_FAST 0
#define WFPT_VARIANT_FAITHFUL 1

/* Library / device housekeeping ------------------------------------------------------------------------------ */
const char* wfpt_last_error(void);
int wfpt_version(void);
/* name[<=name_len], SM count, SM clock (kHz), total memory (bytes) of the current device. */
int wfpt_device_info(char* name, int name_len, int* sm_count, int* clock_khz, int64_t* mem_bytes);
/* Cumulative count of kernels this library has launched in this process (bench.py reports it as gpu_launches). */
int64_t wfpt_launch_count(void);
/* Frees the library's cached device scratch on the current device. */
int wfpt_release_scratch(void);
/* Measures the FP64 issue ceiling on the current device: a register-resident DFMA chain kernel.
 * Returns DFMA instructions per second (lane-level; x2 = FLOP/s) in *dfma_per_s. */
int wfpt_fp64_peak(int iters, double* dfma_per_s, double* ms);

/* Series-term table of the fast variant (host side, no GPU needed).  For a given err the reference's choice of
 * expansion and term count K (src/pdf.pxi:35-60) is piecewise constant in tt; the library finds the piece boundaries
 * by bisection on the reference's expressions with the host libm.  wfpt_ktable copies up to `cap` thresholds (as
 * doubles) and descriptors ((small ? 0x100 : 0) | K) of the pieces starting at them; *desc0 describes the piece below
 * the first threshold; *n == -1 means "not representable" (the kernels then evaluate the formulas directly).
 * wfpt_ktable_decision evaluates the reference decision for one tt (same descriptor encoding, -1 if out of range). */
int wfpt_ktable(double err, int* n, int* kl_max, int* ks_max, int* desc0, double* thr, int* desc, int cap);
int wfpt_ktable_decision(double tt, double err);
int wfpt_ktable_decision_array(const double* tt, int64_t n, double err, int* out);   /* the same for n values */

/* Diagnostic: census of the fast variant's series-term decisions against the reference's.  The reference decides the
 * expansion and K on tt = x'/pow(a,2) (src/pdf.pxi:35-60); the fast kernels on a tt built from a reciprocal.  For every
 * (parameter set, trial inside the support, staged t-node of the integration grid the options select) the kernel forms
 * both and adds to d_counts3[0] the pairs visited, to [1] the pairs whose two tt differ in any bit, to [2] the pairs whose
 * table descriptors differ ("K flips": each would move one density by up to err).  d_counts3 is uint64[3] on the device,
 * zeroed by the caller.  No reference counterpart (the reference has one arithmetic). */
int wfpt_kflip_census(const double* d_x, int64_t n, const double* d_params, int64_t P, double err, int n_st, int n_sz,
                      int use_adaptive, uint64_t* d_counts3, void* stream);

/* Likelihood ------------------------------------------------------------------------------------------------- */

/* P independent wiener_like() evaluations on one RT array:
 *   out[j] = wiener_like(x, *params[j][0:7], err, n_st, n_sz, use_adaptive, simps_err, p_outlier=params[j][7], w_outlier)
 * replaces src/wfpt.pyx:54-76 (one call per row).  No counterpart with P>1 exists in the reference; P=1 is the
 * reference call.  Trials are summed with a deterministic two-stage tree (partials live in library scratch).  Large
 * launches on one RT array run specialised kernels (like_lean.cu): fixed grids with n_sz = 2 in one launch; adaptive
 * integration in two -- the first level, then a refinement kernel over per-set lists of the trials whose recursion goes
 * deeper (library scratch, at most 2 GiB).  Results are independent of the launch geometry to rounding (~1e-15). */
int wfpt_like_batch(const double* d_x, int64_t n, const double* d_params, int64_t P, double err, int n_st, int n_sz,
                    int use_adaptive, double simps_err, double w_outlier, int variant, double* d_out, void* stream);
int wfpt_like_batch_host(const double* h_x, int64_t n, const double* h_params, int64_t P, double err, int n_st,
                         int n_sz, int use_adaptive, double simps_err, double w_outlier, int variant, double* h_out);

/* wfpt_like_batch_host with the RTs reordered by |rt| first (ties in array order, as np.argsort(np.abs(x), kind="stable")):
 * tt = (|rt| - t)/a^2 is then monotone along the thread index, so the lanes of a warp take the same series expansion and K;
 * only the summation order changes (~1e-16 relative).  Up to 32,768 RTs are counted into place on the device (a few
 * microseconds), longer arrays are sorted on the host.  wfpt_sort_rts is the device-to-device form of that reordering
 * (n <= 32,768).  No reference counterpart (a launch-geometry matter). */
int wfpt_like_batch_host_presort(const double* h_x, int64_t n, const double* h_params, int64_t P, double err, int n_st,
                                 int n_sz, int use_adaptive, double simps_err, double w_outlier, int variant, double* h_out);
int wfpt_sort_rts(const double* d_x, int64_t n, double* d_out, void* stream);

/* wfpt_like_batch_host (same arguments, bit-identical results) for the caller HDDM is: a sampler that evaluates
 * wiener_like on the SAME RT array over and over with new parameters (hddm/likelihoods.py:117-121, one array per
 * subject x condition node, hddm/models/base.py:1488-1493; 5-10 calls per stochastic and iteration under SliceStep,
 * hddm/models/hddm_info.py:539-558).  The library keeps a device copy of every array it has been given, keyed by
 * (h_x, n) per device, beside a host copy that every later call is compared with byte for byte: an array changed in
 * place -- or another array at a recycled address -- is noticed and uploaded again.  A hit moves 64 bytes per parameter
 * row.  Least-recently-used entries are dropped beyond wfpt_cache_config's limits (defaults 65,536 arrays, 2 GiB;
 * max_entries == 0 turns the cache off and the call becomes wfpt_like_batch_host).  wfpt_cache_clear frees the current
 * device's entries; wfpt_cache_stats reports them (any pointer may be NULL). */
int wfpt_like_cached_host(const double* h_x, int64_t n, const double* h_params, int64_t P, double err, int n_st,
                          int n_sz, int use_adaptive, double simps_err, double w_outlier, int variant, double* h_out);
int wfpt_cache_config(int64_t max_entries, int64_t max_bytes);
int wfpt_cache_clear(void);
int wfpt_cache_stats(int64_t* entries, int64_t* bytes, int64_t* hits, int64_t* misses, int64_t* stale);

/* The computation of wfpt_like_batch for callers that cannot allocate device memory (a process whose GPU memory is
 * already owned by another pool, a latency-bound caller that cannot afford cudaMalloc's device synchronisation, a
 * context in which allocation calls are not permitted).  It has no counterpart in the reference.
 * The partial sums of the two-stage trial reduction live in a fixed array inside the library image
 * (WFPT_NOALLOC_MAX_SLOTS slots, 192 KiB, held by every GPU that loads the library until process exit) and the
 * function makes no cudaMalloc/cudaFree call.  A call needs P x slices slots, where slices is the library's general
 * trial slicing for (n, P) (one slice once P reaches eight CTAs per SM, otherwise enough slices to get there, at least
 * 256 RTs each); the small-grid re-plan wfpt_like_batch applies on top of that is NOT applied here, so for small grids
 * the two functions cut the trials differently, and large launches of wfpt_like_batch go through kernels that need further
 * library scratch (adaptive integration: lists of trials awaiting refinement) where this one keeps to kernels that need
 * none: the sums of the two
 * agree to rounding (<= 1e-13 relative), not bit for bit.
 * More than WFPT_NOALLOC_MAX_SLOTS slots is refused with an error.  All calls on a device share the array, so calls
 * whose kernels overlap on the GPU race; run them one after another.  The first call on each device looks the array's
 * device address up and should be made before any stream capture; the address is cached per device for the life of the
 * process, so after a cudaDeviceReset (which destroys the context the address belongs to) this entry point must not
 * be used again in that process. */
#define WFPT_NOALLOC_MAX_SLOTS 16384
int wfpt_like_batch_noalloc(const double* d_x, int64_t n, const double* d_params, int64_t P, double err, int n_st,
                            int n_sz, int use_adaptive, double simps_err, double w_outlier, int variant,
                            double* d_out, void* stream);

/* y[i] = full_pdf(x[i], ...)*(1-p_outlier) + w_outlier*p_outlier, optionally log.  Replaces src/wfpt.pyx:32-48
 * (pdf_array) and, with n == 1, p_outlier == 0, the cpdef full_pdf of src/pdf.pxi:104-146. */
int wfpt_pdf_array(const double* d_x, int64_t n, const double* params8, double err, int n_st, int n_sz,
                   int use_adaptive, double simps_err, double w_outlier, int logp, int variant, double* d_y,
                   void* stream);
int wfpt_pdf_array_host(const double* h_x, int64_t n, const double* params8, double err, int n_st, int n_sz,
                        int use_adaptive, double simps_err, double w_outlier, int logp, int variant, double* h_y);

/* Per-trial parameters.  per_trial[k] (k: v, sv, a, z, sz, t, st) is NULL -> scalars[k] is used for every trial,
 * else an array of n doubles.  RTs of +-999 mean "choice only" and contribute log P(upper) / log(1-P(upper)).
 * No early exit: log(0) = -inf flows through the IEEE sum.  Replaces src/wfpt.pyx:357-387 (wiener_like_multi). */
int wfpt_like_multi(const double* d_x, int64_t n, const double* const d_per_trial[7], const double scalars[7],
                    double err, int n_st, int n_sz, int use_adaptive, double simps_err, double p_outlier,
                    double w_outlier, int variant, double* d_out, void* stream);
int wfpt_like_multi_host(const double* h_x, int64_t n, const double* const h_per_trial[7], const double scalars[7],
                         double err, int n_st, int n_sz, int use_adaptive, double simps_err, double p_outlier,
                         double w_outlier, int variant, double* h_out);

/* RL-DDM: Rescorla-Wagner Q recurrence per split_by group (one thread per group, sequential in trial order), then
 * one density per trial with drift (Q_upper - Q_lower) * v; the first trial of each group only updates Q.
 * `order`[n] lists trial indices grouped by ascending split_by value (stable), `starts`[n_groups+1] delimits the
 * groups -- the grouping np.unique + boolean masks produce in src/wfpt.pyx:99,113-116.
 * Replaces src/wfpt.pyx:78-153 (wiener_like_rlddm). */
int wfpt_like_rlddm_host(const double* h_x, const int64_t* h_response, const double* h_feedback,
                         const int64_t* h_order, const int64_t* h_starts, int64_t n_groups, int64_t n, double q,
                         double alpha, double pos_alpha, double v, double sv, double a, double z, double sz, double t,
                         double st, double err, int n_st, int n_sz, int use_adaptive, double simps_err,
                         double p_outlier, double w_outlier, int variant, double* h_out);

/* One RT array per parameter set: set p is evaluated on x[seg[p][0] .. seg[p][1]) (what a hierarchical model's logp
 * sweep is: one wiener_like per (subject x condition) node, hddm/models/base.py:1488-1493, all in one launch; a
 * segment may be named by several sets, e.g. several proposals for one node).  seg is int64[P][2] (start, end);
 * mean_len (average segment length) only steers the launch geometry.  Each out[p] equals wiener_like on that segment
 * (src/wfpt.pyx:54-76).  The _host form copies x[0 .. max end) to the device. */
int wfpt_like_segments(const double* d_x, const int64_t* d_seg, int64_t mean_len, const double* d_params, int64_t P,
                       double err, int n_st, int n_sz, int use_adaptive, double simps_err, double w_outlier, int variant,
                       double* d_out, void* stream);
int wfpt_like_segments_host(const double* h_x, const int64_t* h_seg, const double* h_params, int64_t P, double err,
                            int n_st, int n_sz, int use_adaptive, double simps_err, double w_outlier, int variant,
                            double* h_out);

/* The sampler's form of the above: RTs and segments already resident on the device (uploaded once before sampling),
 * parameters in a host buffer, log-likelihoods back into a host buffer -- one sweep over the nodes of a hierarchical
 * model (hddm/models/hddm_info.py:539-558 evaluates them one wiener_like at a time).  Synchronous. */
int wfpt_like_segments_resident(const double* d_x, const int64_t* d_seg, int64_t mean_len, const double* h_params,
                                int64_t P, double err, int n_st, int n_sz, int use_adaptive, double simps_err,
                                double w_outlier, int variant, double* h_out);

/* The same behind a handle, for callers that hold no device pointers (a sampler driving the library through ctypes
 * only): wfpt_resident_create uploads the packed RT array (each node's RTs |rt|-sorted by the caller, see
 * hddm_b200.wfpt.sort_rts) and the node table seg[n_nodes][2] once; wfpt_resident_logp evaluates one parameter row
 * per node (h_nodes == NULL, P == n_nodes) or per entry of h_nodes (a node may be named several times: several
 * proposals for it), host parameters in, host log-likelihoods out; wfpt_resident_release frees the device memory.
 * wfpt_resident_pointers exposes the device addresses for callers that want to pass device-resident parameters to
 * wfpt_like_segments themselves.  One RT array per likelihood node of a hierarchical model
 * (hddm/models/base.py:1488-1493); no reference counterpart at this level. */
int wfpt_resident_create(const double* h_x, int64_t total, const int64_t* h_seg, int64_t n_nodes, uint64_t* handle);
int wfpt_resident_logp(uint64_t handle, const int64_t* h_nodes, const double* h_params, int64_t P, double err, int n_st,
                       int n_sz, int use_adaptive, double simps_err, double w_outlier, int variant, double* h_out);
int wfpt_resident_pointers(uint64_t handle, const double** d_x, const int64_t** d_seg, int64_t* n_nodes, int64_t* mean_len);
int wfpt_resident_release(uint64_t handle);

/* Replaces src/wfpt.pyx:551-580 (wiener_like_contaminant): trials with cont_x == 0 enter the diffusion likelihood (no
 * outlier mix, -inf on the first zero density), the others contribute n_cont * log(0.5 / (t_max - t_min)). */
int wfpt_like_contaminant_host(const double* h_x, const int32_t* h_cont, int64_t n, double v, double sv, double a, double z,
                               double sz, double t, double st, double t_min, double t_max, double err, int n_st, int n_sz,
                               int use_adaptive, double simps_err, int variant, double* h_out);

/* Replaces src/wfpt.pyx:270-354 (wiener_like_rl): choice-only RL likelihood; grouping arguments as in
 * wfpt_like_rlddm_host. */
int wfpt_like_rl_host(const int64_t* h_response, const double* h_feedback, const int64_t* h_order, const int64_t* h_starts,
                      int64_t n_groups, int64_t n, double q, double alpha, double pos_alpha, double v, double z,
                      double p_outlier, double w_outlier, double* h_out);

/* Host utility for the callers of the two functions above: the grouping `for s in np.unique(split_by): mask = split_by == s`
 * (src/wfpt.pyx:99,109-116) as index lists -- order[n] = trial indices group by group (groups by ascending split_by
 * value, trials in array order inside a group), starts[0..n_groups] = group boundaries in `order`.  starts must hold
 * n + 1 entries.  No device work. */
int wfpt_group_by_split(const int64_t* h_split_by, int64_t n, int64_t* h_order, int64_t* h_starts, int64_t* n_groups);

/* Replaces src/wfpt.pyx:390-433 (wiener_like_multi_rlddm): trials in array order, Q reset where split_by changes,
 * per-trial parameters; per_trial[0..6] = v, sv, a, z, sz, t, st and per_trial[7] = alpha (NULL -> scalars[k]). */
int wfpt_like_multi_rlddm_host(const double* h_x, const int64_t* h_response, const double* h_feedback,
                               const int64_t* h_split_by, int64_t n, double q, const double* const h_per_trial[8],
                               const double scalars[8], double err, int n_st, int n_sz, int use_adaptive,
                               double simps_err, double p_outlier, double w_outlier, int variant, double* h_out);

/* y[i] = CDF of the seven-parameter diffusion model at the signed RT x[i], with the outlier mixture of the wrapper.
 * Replaces hddm.cdfdif.dmat_cdf_array (src/cdfdif_wrapper.pyx:16-53, src/cdfdif.c:36-221; used by Wfpt.cdf,
 * hddm/likelihoods.py:232-233).  Argument validation (ValueError on out-of-support parameters) is the caller's, as
 * in the wrapper. */
int wfpt_dmat_cdf_array_host(const double* h_x, int64_t n, double v, double sv, double a, double z, double sz,
                             double t, double st, double p_outlier, double w_outlier, double* h_y);

/* Simulators ------------------------------------------------------------------------------------------------- */

/* Model ids: which Euler-Maruyama simulator of src/cddm_data_simulation.pyx runs, and the theta column order the
 * reference's dispatch uses for it (graveyard/basic_simulator_old.py:168-480, hddm/model_config.py). */
#define WFPT_SIM_DDM 0            /* ddm :294-375                 theta = v, a, z, t                bounds [0,a], choices 0/1  */
#define WFPT_SIM_FULL_DDM_BASE 1  /* full_ddm_hddm_base :174-284  theta = v, a, z, t, sz, sv, st    bounds [0,a], choices 0/1  */
#define WFPT_SIM_DDM_FLEXBOUND 2  /* ddm_flexbound :454-553, constant bound  theta = v, a, z, t     bounds +-a,  choices -1/1  */
#define WFPT_SIM_ANGLE 3          /* ddm_flexbound, angle bound   theta = v, a, z, t, theta                                    */
#define WFPT_SIM_WEIBULL 4        /* ddm_flexbound, Weibull bound theta = v, a, z, t, alpha, beta                              */
#define WFPT_SIM_FULL_DDM 5       /* full_ddm :747-869            theta = v, a, z, t, sz, sv, st                               */
#define WFPT_SIM_DDM_SDV 6        /* ddm_sdv :876-994             theta = v, a, z, t, sv                                       */
#define WFPT_SIM_ORNSTEIN 7       /* ornstein_uhlenbeck :1119-1219  theta = v, a, z, g, t                                      */
#define WFPT_SIM_LEVY 8           /* levy_flexbound :642-741      theta = v, a, z, alpha, t                                    */
#define WFPT_SIM_GAMMA_DRIFT 9        /* ddm_flex :1001-1113 + gamma_drift        theta = v, a, z, t, shape, scale, c          */
#define WFPT_SIM_GAMMA_DRIFT_ANGLE 10 /* same, angle bound                         theta = v, a, z, t, theta, shape, scale, c   */
#define WFPT_SIM_DS_CONFLICT 11       /* ddm_flex + ds_conflict_drift  theta = a, z, t, tinit, dinit, tslope, dslope, tfixedp, tcoh, dcoh */
#define WFPT_SIM_DS_CONFLICT_ANGLE 12 /* same, angle bound             theta = ... , angle                                     */
#define WFPT_SIM_DDM_COV 13           /* ddm_cov :384-449              theta = v, a, z, t    bounds [0,a], choices -sign(y)      */

/* Parameters per set of a model id, or -1 for an unknown id. */
int wfpt_sim_n_params(int model);

/* n_samples particles for each of n_sets parameter sets (the reference's "n_trials" axis).  theta is float32
 * [n_sets][n_par].  Outputs, each optional (NULL): rts float32 / choices int32, both [n_samples][n_sets] (the
 * reference's (n_samples, n_trials, 1) arrays); hist uint32 [n_sets][nbins][2] RT x choice counts, ADDED to what the
 * buffer holds, with the bin edges of bin_simulator_output (graveyard/basic_simulator_old.py:70-119): nbins edges
 * linspace(0, hist_max_t, nbins) plus +inf, column 0 = lower / first choice, column 1 = upper; steps: total Euler steps
 * (added); traj: path of particle 0 of set 0 (float32[traj_len], caller pre-fills with -999 like the reference).
 * Random numbers: Philox4x32-7 keyed by `seed`, counter = (sample_offset + particle index, set): results do not depend
 * on launch geometry, and ranks that shard particles pass their global offset in `sample_offset`.
 * Replaces the compiled simulators behind hddm.simulators.simulator (hddm/simulators/basic_simulator.py:6-36). */
int wfpt_sim_run(int model, const float* d_theta, int n_par, int n_sets, int64_t n_samples, int64_t sample_offset,
                 float delta_t, float max_t, float s, uint64_t seed, float* d_rts, int32_t* d_choices, uint32_t* d_hist,
                 int nbins, float hist_max_t, uint64_t* d_steps, float* d_traj, int traj_len, void* stream);
/* wfpt_sim_run for one rank's share of the parameter sets: local set k is global set set_offset + k * set_stride, and
 * that global index is what keys the random streams, so histograms and samples of a shard are bit-identical to the same
 * sets' in an unsharded run (ranks sharding sets round-robin pass set_offset = rank, set_stride = world size).
 * wfpt_sim_run is this with (0, 1). */
int wfpt_sim_run_sharded(int model, const float* d_theta, int n_par, int n_sets, int set_offset, int set_stride,
                         int64_t n_samples, int64_t sample_offset, float delta_t, float max_t, float s, uint64_t seed,
                         float* d_rts, int32_t* d_choices, uint32_t* d_hist, int nbins, float hist_max_t,
                         uint64_t* d_steps, float* d_traj, int traj_len, void* stream);
int wfpt_sim_run_host(int model, const float* h_theta, int n_par, int n_sets, int64_t n_samples, int64_t sample_offset,
                      float delta_t, float max_t, float s, uint64_t seed, float* h_rts, int32_t* h_choices,
                      uint32_t* h_hist, int nbins, float hist_max_t, uint64_t* h_steps, float* h_traj, int traj_len);

/* Multi-accumulator simulators: race_model (src/cddm_data_simulation.pyx:1253-1374, lca == 0) and lca (:1380-1518,
 * lca != 0) with n_particles = 2..4 accumulators and a constant (angle == 0) or linearly collapsing (angle != 0) bound.
 * theta is float32 [n_sets][2*n_particles + 5] in the canonical layout v[np], a, z[np], g, b, t, angle (the Python shim
 * builds it from the reference's per-model column orders, graveyard/basic_simulator_old.py:476-760).  choices are
 * accumulator indices (np.argmax); hist is uint32 [n_sets][nbins][n_particles].  Other arguments as in wfpt_sim_run. */
int wfpt_sim_race_run_host(int lca, int n_particles, int angle, const float* h_theta, int n_sets, int64_t n_samples,
                           int64_t sample_offset, float delta_t, float max_t, float s, uint64_t seed, float* h_rts,
                           int32_t* h_choices, uint32_t* h_hist, int nbins, float hist_max_t, uint64_t* h_steps);

/* Two-stage 4-choice simulator ddm_flexbound_par2 (src/cddm_data_simulation.pyx:1667-1791): theta is float32
 * [n_sets][10] = v_h, v_l1, v_l2, a, z_h, z_l1, z_l2, t, b0, b1 with bound_kind 0 constant, 1 angle (b0 = theta),
 * 2 Weibull (b0 = alpha, b1 = beta); choices 0..3; hist is uint32 [n_sets][nbins][4].  Other arguments as in
 * wfpt_sim_run. */
int wfpt_sim_par2_run_host(int bound_kind, const float* h_theta, int n_sets, int64_t n_samples, int64_t sample_offset,
                           float delta_t, float max_t, float s, uint64_t seed, float* h_rts, int32_t* h_choices,
                           uint32_t* h_hist, int nbins, float hist_max_t, uint64_t* h_steps);

/* Siblings of the two-stage model: kind 1 = ddm_flexbound_seq2 (src/cddm_data_simulation.pyx:1523-1661; the
 * low-dimension walk starts when and where the high one ends), 2 = ddm_flexbound_mic2_adj (:1797-1960) and
 * 3 = ddm_flexbound_tradeoff (:1966-2137) (both walks start at time 0, the low one damped by the high walker's
 * position through d).  theta rows carry 13 floats: v_h, v_l1, v_l2, a, z_h, z_l1, z_l2, t, b0, b1, d, b2, b3.
 * bound_kind 0 / 1 / 2 as above, and 3 = the conflict-gamma bound of `tradeoff_conflict_gamma_no_bias`
 * (graveyard/boundary_functions.py:66-97): max(a + b1 * gamma_pdf(t; shape b2, scale b3) + t * (-tan b0), 0).
 * Everything else as wfpt_sim_par2_run_host. */
int wfpt_sim_hier_run_host(int kind, int bound_kind, const float* h_theta, int n_sets, int64_t n_samples,
                           int64_t sample_offset, float delta_t, float max_t, float s, uint64_t seed, float* h_rts,
                           int32_t* h_choices, uint32_t* h_hist, int nbins, float hist_max_t, uint64_t* h_steps);

#ifdef __cplusplus
}
#endif
#endif /* WFPT_B200_H */
