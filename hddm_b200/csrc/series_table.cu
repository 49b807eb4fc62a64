// series_table.cu -- host side: break points of the Navarro-Fuss series selection for one err.
//
// The reference decides, inside every pdf_sv call, which expansion to sum and how many terms
// (src/pdf.pxi:36-54,60): kl and ks are computed from (tt, err) with two logs and four square
// roots, then `ks < kl` picks the branch and ceil() the term count K.  The pair (branch, K) is a
// piecewise-constant function of tt for a fixed err.  Here the host evaluates the reference's
// formulas (same operation order, glibc libm, no FMA contraction on x86-64) on a scan of the
// whole positive double range and bisects every change down to adjacent doubles, so the device
// can replace ~200 FP64 instructions per time node by a handful of compares -- and the decisions
// are the host libm's, i.e. the reference's own, not the device libm's.
#include <cmath>
#include <cstring>
#include <map>
#include <mutex>

#include "kernels.cuh"

namespace hddm {

namespace {

// (branch, K) code of src/pdf.pxi:36-54,60: +K small-time, -K large-time, 0 = not representable.
int decide(double tt, double err)
{
    double kl, ks;
    if (M_PI * tt * err < 1) {
        kl = sqrt(-2 * log(M_PI * tt * err) / (pow(M_PI, 2.0) * tt));
        const double fl = 1. / (M_PI * sqrt(tt));
        kl = kl < fl ? fl : kl;
    } else {
        kl = 1. / (M_PI * sqrt(tt));
    }
    if (2 * sqrt(2 * M_PI * tt) * err < 1) {
        ks = 2 + sqrt(-2 * tt * log(2 * sqrt(2 * M_PI * tt) * err));
        const double fl = sqrt(tt) + 1;
        ks = ks < fl ? fl : ks;
    } else {
        ks = 2;
    }
    if (ks < kl) {
        const double K = ceil(ks);
        if (!(K >= 1 && K <= 30000)) return 0;
        return (int)K;
    }
    const double K = ceil(kl);
    if (!(K >= 0 && K <= 30000)) return 0;  // K == 0 is possible for huge tt: empty sum, p = 0
    return K == 0 ? -30001 : -(int)K;
}

double from_bits(uint64_t b)
{
    double d;
    memcpy(&d, &b, sizeof(d));
    return d;
}

struct Builder {
    double err;
    SeriesTable t;
    int n_regions = 0;
    bool overflow = false;

    void emit(double thr, int code)
    {
        if (n_regions >= kMaxRegions || code == 0) {
            overflow = true;
            return;
        }
        if (n_regions > 0) t.thr[n_regions - 1] = thr;
        t.code[n_regions++] = (short)code;
    }
    // code(lo) == clo is known; find every change in (lo, hi]
    void refine(uint64_t lo, int clo, uint64_t hi, int chi)
    {
        if (overflow || clo == chi) return;
        if (hi - lo == 1) {
            emit(from_bits(hi), chi);
            return;
        }
        const uint64_t mid = lo + (hi - lo) / 2;
        const int cmid = decide(from_bits(mid), err);
        refine(lo, clo, mid, cmid);
        refine(mid, cmid, hi, chi);
    }
};

SeriesTable build(double err)
{
    Builder b;
    b.err = err;
    memset(&b.t, 0, sizeof(b.t));
    b.t.err = err;
    // uniform steps in the bit pattern are logarithmic steps in tt: 64 points per binade from the
    // smallest subnormal to DBL_MAX.
    const uint64_t first = 1, last = 0x7fefffffffffffffULL, step = 1ULL << 46;
    uint64_t lo = first;
    int clo = decide(from_bits(lo), err);
    b.emit(0.0, clo);
    while (lo < last && !b.overflow) {
        const uint64_t hi = (last - lo > step) ? lo + step : last;
        const int chi = decide(from_bits(hi), err);
        if (clo != chi) {
            b.refine(lo, clo, hi, chi);
        } else {  // same code at both ends: one interior probe guards against a double flip
            const uint64_t mid = lo + (hi - lo) / 2;
            const int cmid = decide(from_bits(mid), err);
            if (cmid != clo) {
                b.refine(lo, clo, mid, cmid);
                b.refine(mid, cmid, hi, chi);
            }
        }
        lo = hi;
        clo = chi;
    }
    SeriesTable t = b.t;
    if (b.overflow || !(err == err)) {
        t.n_thr = -1;
        return t;
    }
    t.n_thr = b.n_regions - 1;
    for (int r = 0; r < b.n_regions; ++r) {
        int c = t.code[r];
        if (c == -30001) {  // K == 0: keep as an empty large-time sum
            t.code[r] = 0;
            continue;
        }
        if (c > 0 && c > t.max_small_k) t.max_small_k = c;
        if (c < 0 && -c > t.max_large_k) t.max_large_k = -c;
    }
    return t;
}

std::mutex g_mu;
std::map<uint64_t, SeriesTable> g_cache;

}  // namespace

const SeriesTable& series_table_for(double err)
{
    uint64_t key;
    memcpy(&key, &err, sizeof(key));
    std::lock_guard<std::mutex> lock(g_mu);
    auto it = g_cache.find(key);
    if (it == g_cache.end()) it = g_cache.emplace(key, build(err)).first;
    return it->second;
}

}  // namespace hddm
