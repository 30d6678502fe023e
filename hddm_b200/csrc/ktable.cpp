// Host-side construction of the series-term table (see ktable.h).
#include "ktable.h"

#include <cmath>
#include <cstring>
#include <map>
#include <mutex>
#include <vector>

namespace wfpt {

namespace {

inline double max_std(double a, double b) { return (a < b) ? b : a; }

inline long long bits_of(double x)
{
    long long b;
    std::memcpy(&b, &x, sizeof(b));
    return b;
}
inline double from_bits(long long b)
{
    double x;
    std::memcpy(&x, &b, sizeof(x));
    return x;
}

}  // namespace

// /root/reference/src/pdf.pxi:35-60, same expression order; volatile keeps the compiler from re-associating.
int ktable_reference_decision(double tt, double err)
{
    const double PI = 3.14159265358979323846;
    volatile double kl, ks;
    if (PI * tt * err < 1) {
        kl = std::sqrt(-2 * std::log(PI * tt * err) / ((PI * PI) * tt));
        kl = max_std(kl, 1. / (PI * std::sqrt(tt)));
    } else {
        kl = 1. / (PI * std::sqrt(tt));
    }
    if (2 * std::sqrt(2 * PI * tt) * err < 1) {
        ks = 2 + std::sqrt(-2 * tt * std::log(2 * std::sqrt(2 * PI * tt) * err));
        ks = max_std(ks, std::sqrt(tt) + 1);
    } else {
        ks = 2;
    }
    const bool small = ks < kl;
    const double kc = std::ceil(small ? ks : kl);
    if (!(kc >= 0 && kc <= 255)) return -1;  // not representable in the table
    return (small ? 0x100 : 0) | (int)kc;
}

static void refine(double lo, double hi, int dlo, int dhi, double err, std::vector<std::pair<double, int>>& out, int depth)
{
    // invariant: decision(lo) == dlo != dhi == decision(hi), lo < hi.  Finds every change point in (lo, hi].
    long long a = bits_of(lo), b = bits_of(hi);
    while (b - a > 1) {
        const long long m = a + (b - a) / 2;
        const int dm = ktable_reference_decision(from_bits(m), err);
        if (dm == dlo) {
            a = m;
        } else if (dm == dhi || depth > 40) {
            b = m;
        } else {
            // a third value in between: split
            refine(from_bits(a), from_bits(m), dlo, dm, err, out, depth + 1);
            refine(from_bits(m), from_bits(b), dm, dhi, err, out, depth + 1);
            return;
        }
    }
    out.emplace_back(from_bits(b), ktable_reference_decision(from_bits(b), err));
}

static KTable build_ktable(double err)
{
    KTable kt;
    std::memset(&kt, 0, sizeof(kt));
    kt.n = -1;
    if (!(err > 0) || !std::isfinite(err)) return kt;

    // scan positive doubles from the smallest subnormal to DBL_MAX on a 2^(1/32) grid
    std::vector<std::pair<double, int>> changes;
    double prev = from_bits(1);
    int dprev = ktable_reference_decision(prev, err);
    if (dprev < 0) return kt;
    const int first = dprev;
    const double step = std::exp2(1.0 / 32.0);
    double cur = prev;
    for (;;) {
        double nxt = cur * step;
        if (!(nxt > cur)) nxt = std::nextafter(cur, INFINITY);
        if (cur < 1e-300) nxt = cur * 2 > cur ? cur * 2 : nxt;  // coarse steps through the subnormals
        if (!std::isfinite(nxt)) nxt = 1.7976931348623157e308;
        const int d = ktable_reference_decision(nxt, err);
        if (d < 0) return kt;
        if (d != dprev) refine(cur, nxt, dprev, d, err, changes, 0);
        dprev = d;
        if (nxt >= 1.7976931348623157e308) break;
        cur = nxt;
    }
    if ((int)changes.size() > kMaxPieces) return kt;
    kt.desc0 = first;
    int kl_max = 0, ks_max = 0;
    auto track = [&](int d) {
        if (d & 0x100) ks_max = std::max(ks_max, d & 0xff);
        else kl_max = std::max(kl_max, d & 0xff);
    };
    track(first);
    for (size_t i = 0; i < changes.size(); i++) {
        if (changes[i].second < 0) return kt;
        if (i > 0 && !(changes[i].first > changes[i - 1].first)) return kt;
        kt.thr[i] = bits_of(changes[i].first);
        kt.desc[i] = changes[i].second;
        track(changes[i].second);
    }
    kt.kl_max = kl_max;
    kt.ks_max = ks_max;
    kt.n = (int)changes.size();
    return kt;
}

KTable get_ktable(double err)
{
    static std::mutex mu;
    static std::map<long long, KTable> cache;
    std::lock_guard<std::mutex> lk(mu);
    const long long key = bits_of(err);
    auto it = cache.find(key);
    if (it == cache.end()) {
        if (cache.size() > 256) cache.clear();
        it = cache.emplace(key, build_ktable(err)).first;
    }
    return it->second;
}

}  // namespace wfpt

// C hooks for the tests (declared in include/wfpt_b200.h)
extern "C" int wfpt_ktable(double err, int* n, int* kl_max, int* ks_max, int* desc0, double* thr, int* desc, int cap)
{
    const wfpt::KTable kt = wfpt::get_ktable(err);
    if (n) *n = kt.n;
    if (kl_max) *kl_max = kt.kl_max;
    if (ks_max) *ks_max = kt.ks_max;
    if (desc0) *desc0 = kt.desc0;
    for (int i = 0; i < kt.n && i < cap; i++) {
        if (thr) std::memcpy(&thr[i], &kt.thr[i], sizeof(double));
        if (desc) desc[i] = kt.desc[i];
    }
    return 0;
}

extern "C" int wfpt_ktable_decision(double tt, double err) { return wfpt::ktable_reference_decision(tt, err); }

extern "C" int wfpt_ktable_decision_array(const double* tt, int64_t n, double err, int* out)
{
    for (int64_t i = 0; i < n; i++) out[i] = wfpt::ktable_reference_decision(tt[i], err);
    return 0;
}
