// Series-term table: the reference fixes, per evaluation of f(tt|0,1,w), the expansion (small- or large-time) and
// the number of terms K from (tt, err) alone (/root/reference/src/pdf.pxi:35-60).  For a given err that decision is
// a piecewise-constant function of tt with a handful of pieces.  build_ktable() finds the piece boundaries to the
// exact double by bisection on the reference's own expressions, evaluated with the host's libm (the same glibc the
// reference's extension links), so the device can replace 2 log + 6 sqrt + 2 div per node by a few integer compares
// and still take exactly the reference's K (the device's own log/sqrt would differ from glibc in the last ulp).
#pragma once
#include <stdint.h>

namespace wfpt {

constexpr int kMaxPieces = 32;   // thresholds the device table can hold; more pieces -> table unusable (n = -1)
constexpr int kMaxKLarge = 8;    // large-time terms the fast evaluator keeps in registers

struct KTable {
    int n;                        // number of thresholds (pieces = n + 1); -1: not representable, use the formulas
    int kl_max;                   // largest K among large-time pieces
    int ks_max;                   // largest K among small-time pieces
    int desc0;                    // descriptor of the first piece (tt below thr[0])
    long long thr[kMaxPieces];    // bit patterns of the ascending thresholds: piece i+1 starts at tt >= thr[i]
    int desc[kMaxPieces];         // descriptor of the piece starting at thr[i]:  (small ? 0x100 : 0) | K
};

// Host only.  Returns a copy of the table for `err` (cached per distinct err value; copied under the cache's lock, so a
// later eviction cannot touch what a caller holds).
KTable get_ktable(double err);

// The reference decision itself (host libm), exposed for tests: returns (small ? 0x100 : 0) | K.
int ktable_reference_decision(double tt, double err);

}  // namespace wfpt
