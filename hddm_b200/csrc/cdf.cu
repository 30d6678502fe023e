// CDF of the seven-parameter diffusion model (Tuerlinckx 2004) on the GPU: the computation behind
// hddm.cdfdif.dmat_cdf_array (/root/reference/src/cdfdif_wrapper.pyx:16-53, which calls cdfdif() of
// /root/reference/src/cdfdif.c:36-221), used by Wfpt.cdf for chi-square / G-square quantile optimisation
// (/root/reference/hddm/likelihoods.py:232-233).
//
// One thread per RT.  What is computed is the reference's scheme: drift variability by 6-point Gauss-Hermite, starting
// point variability by 6-point Gauss-Legendre, non-decision-time variability analytically (three regimes: RT before,
// inside, after the window of possible non-decision times), eigenfunction series summed until three successive partial
// sums agree to 1e-29 or 5,000 terms.  The file is compiled with --fmad=false and keeps the reference's order of
// operations inside each term so that the data-dependent stopping point and the heavy cancellation see the same
// roundings (up to libm differences).
//
// Conditioning note: the wrapper replaces sv = 0, sz = 0, st = 0 by 1e-10-sized values and the series are then divided
// by sz*st; the reference's own result carries ~1e-6 relative noise there, and so does any evaluation with a different
// libm.  tests/test_gpu_cdf.py compares at 1e-9 for sz, st away from 0 and at 1e-5 absolute otherwise.
#include "../../include/wfpt_b200.h"
#include "common.cuh"

namespace wfpt {
namespace cdf {

constexpr double kPiC = 3.1415926535897932384626433832795028841971693993751;
constexpr int kTermCap = 5000;
constexpr double kAgree = 1e-29;   // partial sums closer than this count as converged
constexpr double kTinyDrift = 1e-7;
constexpr double kMinDecision = 0.001;

__constant__ double kHermiteX[6] = {-2.3506049736744922818, -1.3358490740136970132, -.43607741192761650950,
                                    .43607741192761650950, 1.3358490740136970132, 2.3506049736744922818};
__constant__ double kHermiteW[6] = {.45300099055088421593e-2, .15706732032114842368, .72462959522439207571,
                                    .72462959522439207571, .15706732032114842368, .45300099055088421593e-2};
__constant__ double kLegendreX[6] = {-.93246951420315193904, -.66120938646626381541, -.23861918608319693247,
                                     .23861918608319712676, .66120938646626459256, .93246951420315160597};
__constant__ double kLegendreW[6] = {.17132449237917049545, .36076157304813916138, .46791393457269092604,
                                     .46791393457269092604, .36076157304813843973, .17132449237917132812};

// Model in the units of cdfdif.c (the wrapper divides a, sv, v by 10 and scales z, sz by a/10).
struct Model {
    double a, ter, eta, z, sz, st, nu;
    double drift[6], wdrift[6];  // Gauss-Hermite nodes / weights mapped onto N(nu, eta)   (cdfdif.c:71-77)
    double start[6];             // Gauss-Legendre nodes mapped onto [z - sz/2, z + sz/2]  (:78-79)
};

__device__ __forceinline__ void map_quadrature(Model& m)
{
    for (int i = 0; i < 6; i++) {
        m.drift[i] = 1.41421356237309505 * kHermiteX[i] * m.eta + m.nu;
        m.wdrift[i] = kHermiteW[i] / 1.772453850905515882;
        m.start[i] = (.5 * m.sz * kLegendreX[i]) + m.z;
    }
}

// Three-term history of partial sums with the reference's stopping rule (:126-128 and twice more).
struct Partial {
    double h0 = 0, h1 = 0, h2 = 0;
    __device__ __forceinline__ void shift() { h0 = h1; h1 = h2; }
    __device__ __forceinline__ bool settled() const
    {
        return (fabs(h0 - h1) < kAgree) && (fabs(h1 - h2) < kAgree) && (h2 > 0);
    }
};

// P(absorption at the boundary coded by the wrapper), integrated over drift and starting point (:82-100).
__device__ double absorb_prob(const Model& m)
{
    double over_z = 0;
    for (int i = 0; i < 6; i++) {
        double over_drift = 0;
        for (int k = 0; k < 6; k++) {
            if (fabs(m.drift[k]) > kTinyDrift)
                over_drift += (exp(-200 * m.start[i] * m.drift[k]) - 1) / (exp(-200 * m.a * m.drift[k]) - 1) * m.wdrift[k];
            else
                over_drift += m.start[i] / m.a * m.wdrift[k];
        }
        over_z += over_drift * kLegendreW[i] / 2;
    }
    return over_z;
}

// RT beyond the latest possible non-decision time: one series over the eigenvalue index, each term integrated over the
// drift nodes, starting-point and non-decision-time windows in closed form (:106-132).
__device__ double series_after_window(const Model& m, int x, double t, double t_lo, double t_hi, double z_hi, double z_lo)
{
    const double a2 = m.a * m.a;
    Partial s;
    for (int v = 0; v < kTermCap; v++) {
        s.shift();
        double term = 0;
        const double freq = kPiC * v / m.a;
        for (int k = 0; k < 6; k++) {
            const double lam = (100 * m.drift[k] * m.drift[k] + (kPiC * kPiC) * (v * v) / (100 * a2));
            const double up = exp((2 * x - 1) * z_hi * m.drift[k] * 100 - 3 * log(lam) + log(m.wdrift[k]) - 2 * log(100.));
            const double lo = exp((2 * x - 1) * z_lo * m.drift[k] * 100 - 3 * log(lam) + log(m.wdrift[k]) - 2 * log(100.));
            const double edge = up * ((2 * x - 1) * m.drift[k] * sin(freq * z_hi) * 100 - freq * cos(freq * z_hi)) -
                                lo * ((2 * x - 1) * m.drift[k] * sin(freq * z_lo) * 100 - freq * cos(freq * z_lo));
            const double decay = exp((-.5 * lam * (t - t_hi)) + log(1 - exp(-.5 * lam * (t_hi - t_lo))));
            term += edge * decay;
        }
        s.h2 = s.h1 + v * term;
        if (s.settled()) break;
    }
    return s.h2;
}

// RT inside the window, non-negligible drift node k: per starting-point node a closed-form part plus a series (:141-169).
__device__ double window_term_drift(const Model& m, int x, int k, double t, double t_lo)
{
    const double a2 = m.a * m.a, g = m.drift[k];
    double over_z = 0;
    for (int i = 0; i < 6; i++) {
        const double zz = (m.a - m.start[i]) * x + m.start[i] * (1 - x);
        const double closed = -((m.a * a2) / ((1 - 2 * x) * g * kPiC * .01)) * sinh(zz * (1 - 2 * x) * g / .01) /
                                  (sinh((1 - 2 * x) * g * m.a / .01) * sinh((1 - 2 * x) * g * m.a / .01)) +
                              (zz * a2) / ((1 - 2 * x) * g * kPiC * .01) * cosh((m.a - zz) * (1 - 2 * x) * g / .01) /
                                  sinh((1 - 2 * x) * g * m.a / .01);
        Partial s;
        for (int v = 0; v < kTermCap; v++) {
            s.shift();
            const double freq = kPiC * v / m.a;
            const double lam = (g * g * 100 + (kPiC * v) * (kPiC * v) / (a2 * 100));
            s.h2 = s.h1 + v * sin(freq * zz) * exp(-.5 * lam * (t - t_lo) - 2 * log(lam));
            if (s.settled()) break;
        }
        over_z += .5 * kLegendreW[i] * (closed - 4 * s.h2) * (kPiC / 100) / (a2 * m.st) * exp((2 * x - 1) * zz * g * 100);
    }
    return over_z;
}

// RT inside the window, drift node numerically zero: polynomial part plus a cosine series (:171-192).
__device__ double window_term_nodrift(const Model& m, double t, double t_lo, double z_hi, double z_lo)
{
    const double a2 = m.a * m.a;
    const double pu = -(z_hi * z_hi) / (12 * a2) + (z_hi * z_hi * z_hi) / (12 * m.a * a2) - (z_hi * z_hi * z_hi * z_hi) / (48 * a2 * a2);
    const double pl = -(z_lo * z_lo) / (12 * a2) + (z_lo * z_lo * z_lo) / (12 * m.a * a2) - (z_lo * z_lo * z_lo * z_lo) / (48 * a2 * a2);
    Partial s;
    for (int v = 1; v < kTermCap; v++) {
        s.shift();
        const double freq = kPiC * v / m.a;
        const double lam = (kPiC * v) * (kPiC * v) / (a2 * 100);
        s.h2 = s.h1 + 1 / (kPiC * kPiC * kPiC * kPiC * v * v * v * v) * (cos(freq * z_lo) - cos(freq * z_hi)) *
                          exp(-.5 * lam * (t - t_lo));
        if (s.settled()) break;
    }
    return 400 * a2 * m.a * (pl - pu - s.h2) / (m.st * m.sz);
}

// F(t, x): defective CDF of absorption at boundary x by time t; *p_abs = P(absorption at the wrapper's boundary code).
__device__ double defective_cdf(const Model& m, double t, int x, double* p_abs)
{
    const double z_hi = (1 - x) * m.z + x * (m.a - m.z) + m.sz / 2;
    const double z_lo = (1 - x) * m.z + x * (m.a - m.z) - m.sz / 2;
    const double t_lo = m.ter - m.st / 2;
    const double prob = absorb_prob(m);
    *p_abs = prob;
    double F = 0;
    if (t - m.ter + m.st / 2 > kMinDecision) {
        const double t_hi = t < m.ter + m.st / 2 ? t : m.ter + m.st / 2;
        const double p1 = prob * (t_hi - t_lo) / m.st;
        const double p0 = (1 - prob) * (t_hi - t_lo) / m.st;
        const double a2 = m.a * m.a;
        if (t > m.ter + m.st / 2) {
            const double series = series_after_window(m, x, t, t_lo, t_hi, z_hi, z_lo);
            F = (p0 * (1 - x) + p1 * x) - series * 4 * kPiC / (a2 * m.sz * m.st);
        } else {
            double over_drift = 0;
            for (int k = 0; k < 6; k++) {
                const double part = (fabs(m.drift[k]) > kTinyDrift) ? window_term_drift(m, x, k, t, t_lo)
                                                                    : window_term_nodrift(m, t, t_lo, z_hi, z_lo);
                over_drift += part * m.wdrift[k];
            }
            F = (p0 * (1 - x) + p1 * x) - over_drift;
        }
    }
    return F > kAgree ? F : 0;
}

// cdfdif_wrapper.pyx:36-51: unit conversion, signed-RT convention, outlier mixture
__global__ void dmat_cdf_kernel(const double* __restrict__ rt, int64_t n, double v, double sv, double a, double z, double sz,
                                double t, double st, double p_outlier, double w_outlier, double* __restrict__ out)
{
    const int64_t i = (int64_t)blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double nudge = 1e-10;
    Model m;
    m.a = a / 10.;
    m.ter = t;
    m.eta = sv / 10. + nudge;
    m.z = z * (a / 10.);
    m.sz = sz * (a / 10.) + nudge;
    m.st = st + nudge;
    m.nu = v / 10.;
    map_quadrature(m);
    const double x = rt[i];
    double p_abs;
    const double F = defective_cdf(m, fabs(x), (int)(x > 0), &p_abs);
    const double side = (double)((x > 0) - (x < 0));
    const double y = (1 - p_abs) + side * F;
    out[i] = y * (1 - p_outlier) + (x + (1. / (2 * w_outlier))) * w_outlier * p_outlier;
}

}  // namespace cdf
}  // namespace wfpt

using namespace wfpt;

extern "C" int wfpt_dmat_cdf_array_host(const double* h_x, int64_t n, double v, double sv, double a, double z, double sz,
                                        double t, double st, double p_outlier, double w_outlier, double* h_y)
{
    if (n < 0) return api_fail("wfpt_dmat_cdf_array_host: negative size");
    if (n == 0) return 0;
    if (!h_x || !h_y) return api_fail("wfpt_dmat_cdf_array_host: null pointer argument");
    HostCall guard_;
    // staging in the library's grow-only scratch (as every other host entry point): no allocation after the first call
    void *px = nullptr, *py = nullptr;
    if (api_scratch_pair(sizeof(double) * (size_t)n, sizeof(double) * (size_t)n, &px, &py)) return 1;
    double *dx = (double*)px, *dy = (double*)py;
    cudaError_t e = cudaMemcpyAsync(dx, h_x, sizeof(double) * (size_t)n, cudaMemcpyHostToDevice, 0);
    if (e == cudaSuccess) {
        cdf::dmat_cdf_kernel<<<(unsigned)((n + 63) / 64), 64>>>(dx, n, v, sv, a, z, sz, t, st, p_outlier, w_outlier, dy);
        e = cudaGetLastError();
        count_launch(1);
    }
    if (e == cudaSuccess) e = cudaMemcpyAsync(h_y, dy, sizeof(double) * (size_t)n, cudaMemcpyDeviceToHost, 0);
    if (e == cudaSuccess) e = cudaStreamSynchronize(0);
    if (e != cudaSuccess) return api_fail(cudaGetErrorString(e));
    return 0;
}
