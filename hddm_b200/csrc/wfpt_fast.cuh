// wfpt_fast.cuh -- the hot path proper: full_pdf for one (trial, parameter set) when the start
// point (z, sz) is uniform over the CTA's parameter set and integration is adaptive.
//
// Design (not the reference's call graph): the reference evaluates pdf_sv(x - t', v, sv, a, w)
// as a black box at every node of a Simpson tensor grid (src/integrate.pxi:72-206), i.e. per
// node 2 log + 4 sqrt for the series selection, K exp (+ K sin), and log/exp/sqrt/3 div for the
// sv marginal (src/pdf.pxi:28-102).  Here the work is split by what it depends on:
//   * per parameter set and boundary (shared memory, built once per CTA): the start-point nodes
//     w_m = zlo + m*sz/16 of the first three Simpson levels (built by the reference's own midpoint
//     recursion, so they are bit-identical to its abscissae) and sin(k*pi*w_m), k = 1..KMAX;
//   * per decision-time node x_t = x - t': normalised time, the (branch, K) pair from the
//     host-built SeriesTable, q = exp(-pi^2 tt/2) (large-time terms are k*q^(k^2)) or 0.5/tt
//     (small-time), and the sv-marginal factors 1/(2 sv^2 x_t + 2), 1/(sqrt(sv^2 x_t+1) a^2);
//   * per (time node, start-point node): K FMAs (large-time) or K exp (small-time) + one exp.
// The first inner level (5 start points, always needed: src/integrate.pxi:131-133,99-100) is
// evaluated as five independent instruction streams; deeper levels, which the reference's stop
// test rarely asks for, go through a generic node evaluator and an explicit-stack walk.
// DECISIONS (support test, branch, K, every Simpson stop test, the recursion tree) are the
// reference's; values differ from it only by libm / FMA / reassociation rounding (~1e-15).
#pragma once
#include "wfpt_device.cuh"

namespace hddm {

constexpr int kWNodes = 17;    // start-point nodes m/16, m = 0..16
constexpr int kPosShift = 22;  // walk positions are in units of width / 2^26; table node = pos >> 22
constexpr int kPosWidth = 1 << 26;

template <int KMAX>
struct WTable {
    double w[kWNodes];
    double s[KMAX][kWNodes];  // s[k-1][m] = sin((k*pi) * w[m])   (src/pdf.pxi:62)
};

// A cooperating group of threads for the table builders: the whole CTA (barrier = __syncthreads) or one warp
// (barrier = __syncwarp, so that several sets are built side by side in one CTA).
struct BlockGroup {
    __device__ __forceinline__ int tid() const { return threadIdx.x; }
    __device__ __forceinline__ int size() const { return blockDim.x; }
    __device__ __forceinline__ void sync() const { __syncthreads(); }
};
struct WarpGroup {
    __device__ __forceinline__ int tid() const { return threadIdx.x & 31; }
    __device__ __forceinline__ int size() const { return 32; }
    __device__ __forceinline__ void sync() const { __syncwarp(); }
};
// Virtual thread ids of a lane: a CTA thread has one (its own); lane l of a warp takes (l - 5 j) mod 32 + 32 j for
// j = 0..7, a rotation per group of 32 so that the few heavy ids of build_gridset (64; 80, 81; 0..4 and 32..39 of the
// last phase) land on different lanes.
__device__ __forceinline__ int virtual_tid(const BlockGroup&, int tid, int) { return tid; }
__device__ __forceinline__ int virtual_tid(const WarpGroup&, int lane, int j) { return ((lane - 5 * j) & 31) + 32 * j; }
__device__ __forceinline__ int virtual_rounds(const BlockGroup&) { return 1; }
__device__ __forceinline__ int virtual_rounds(const WarpGroup&) { return 8; }

// Cooperative build of the two tables (index 0: lower boundary, start point z; index 1:
// upper boundary, start point 1 - z, src/pdf.pxi:116-118) for one parameter set.
template <int KMAX, class Group = BlockGroup>
__device__ __forceinline__ void build_wtables(WTable<KMAX>* T, double z, double sz_raw, Group grp = Group())
{
    if (grp.tid() < 2) {
        double* w = T[grp.tid()].w;
        const double zz = grp.tid() ? 1. - z : z;
        const double sz = sz_raw < 1e-3 ? 0.0 : sz_raw;  // src/pdf.pxi:124-125
        w[0] = zz - sz / 2.;
        w[16] = zz + sz / 2.;
        for (int step = 8; step >= 1; step >>= 1)  // src/integrate.pxi:83-85: (lb+ub)/2. recursively
            for (int m = step; m < 16; m += 2 * step) w[m] = (w[m - step] + w[m + step]) / 2.;
    }
    grp.sync();
    for (int item = grp.tid(); item < 2 * KMAX * kWNodes; item += grp.size()) {
        const int b = item / (KMAX * kWNodes), r = item % (KMAX * kWNodes);
        const int k = r / kWNodes + 1, m = r % kWNodes;
        T[b].s[k - 1][m] = sin(((double)k * kPi) * T[b].w[m]);
    }
    grp.sync();
}

// Per-trial constants after the boundary flip.
struct TrialCtx {
    double a_sq, inv_a_sq, sv_sq, v_sq;
    double asv, av2;  // a*sv and 2*a*v: numerator (asv*w)^2 - av2*w of src/pdf.pxi:102
    double c5[5];     // that numerator at the five first-level start points
    double inv_zt;    // 1/(ub_z - lb_z) when integrating over z (src/integrate.pxi:129,134), else 1
    const SeriesTable* tab;
    bool sv_zero;
};

// Everything pdf_sv needs that depends on the decision time x_t only.
struct TimeNode {
    double arg;      // small-time: 0.5/tt      large-time: q = exp(-pi^2 tt / 2)
    double scale;    // series post-factor * 1/(sqrt(sv^2 x_t + 1) a^2) * inv_zt
    double inv_den;  // 1/(2 sv^2 x_t + 2)
    double vx;       // v^2 x_t
    int code;        // +K small-time, -K large-time
};

__device__ __forceinline__ TimeNode make_time_node(const TrialCtx& c, double x_t)
{
    TimeNode n;
    const double tt = x_t / c.a_sq;  // src/pdf.pxi:98 (exact division: it feeds the decision)
    n.code = series_code(*c.tab, tt);
    const double sx = c.sv_sq * x_t;
    n.inv_den = 1.0 / (2.0 * sx + 2.0);
    n.vx = c.v_sq * x_t;
    const double out = rsqrt(sx + 1.0) * c.inv_a_sq * c.inv_zt;
    if (n.code > 0) {  // src/pdf.pxi:55-57
        n.arg = 0.5 / tt;
        n.scale = rsqrt((2.0 * kPi) * (tt * tt * tt)) * out;
    } else {  // :61-63
        n.arg = exp(-kPiSq * tt / 2.0);
        n.scale = kPi * out;
    }
    return n;
}

// pdf_sv / ZT at ONE start point w (generic; `m` >= 0 names a table node, else sin is computed).
template <int KMAX>
__device__ __noinline__ double node_value(const TrialCtx& c, const WTable<KMAX>& T, const TimeNode& n,
                                          double w, int m)
{
    double p = 0.0;
    if (n.code > 0) {
        const int K = n.code;
        for (int k = -((K - 1) / 2); k <= K / 2; ++k) {  // src/pdf.pxi:52-56
            const double u = w + (double)(2 * k);
            p += u * exp(-(u * u) * n.arg);
        }
    } else {
        const int K = -n.code;
        const double q2 = n.arg * n.arg;
        double odd = n.arg, pw = n.arg;  // q^(2k-1), q^(k^2)
        for (int k = 1; k <= K; ++k) {   // src/pdf.pxi:60-62
            if (k > 1) {
                odd *= q2;
                pw *= odd;
            }
            const double kd = (double)k;
            const double s = (m >= 0 && k <= KMAX) ? T.s[k - 1][m] : sin((kd * kPi) * w);
            p += (kd * pw) * s;
        }
    }
    if (!c.sv_zero && p < 0.0) return __longlong_as_double(0x7ff8000000000000LL);  // log(p<0)
    const double t1 = c.asv * w;
    const double e = ((t1 * t1 - c.av2 * w) - n.vx) * n.inv_den;
    return (p * n.scale) * exp(e);  // one rounding if the product lands in the subnormal range
}

// The five first-level start points (table nodes 0, 4, 8, 12, 16) of one time node.
template <int KMAX>
__device__ __forceinline__ void first_level(const TrialCtx& c, const WTable<KMAX>& T, const TimeNode& n,
                                            double f[5])
{
    double p[5] = {0.0, 0.0, 0.0, 0.0, 0.0};
    if (n.code > 0) {
        const int K = n.code;
        for (int k = -((K - 1) / 2); k <= K / 2; ++k) {
            const double two_k = (double)(2 * k);
#pragma unroll
            for (int i = 0; i < 5; ++i) {
                const double u = T.w[4 * i] + two_k;
                p[i] += u * exp(-(u * u) * n.arg);
            }
        }
    } else {
        const int K = -n.code;
        const double q2 = n.arg * n.arg;
        double odd = n.arg, pw = n.arg;
#pragma unroll
        for (int k = 1; k <= KMAX; ++k) {
            if (k <= K) {
                if (k > 1) {
                    odd *= q2;
                    pw *= odd;
                }
                const double e = (double)k * pw;
#pragma unroll
                for (int i = 0; i < 5; ++i) p[i] += e * T.s[k - 1][4 * i];
            }
        }
        for (int k = KMAX + 1; k <= K; ++k) {  // never taken when the host picked KMAX >= max K
            odd *= q2;
            pw *= odd;
            const double kd = (double)k;
#pragma unroll
            for (int i = 0; i < 5; ++i) p[i] += (kd * pw) * sin((kd * kPi) * T.w[4 * i]);
        }
    }
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        const double e = (c.c5[i] - n.vx) * n.inv_den;
        double v = (p[i] * n.scale) * exp(e);
        if (!c.sv_zero && p[i] < 0.0) v = __longlong_as_double(0x7ff8000000000000LL);
        f[i] = v;
    }
}

// One pending interval of the adaptive Simpson walk (src/integrate.pxi:72-112, :143-178).
struct Span {
    double lo, hi, coarse, f_lo, f_hi, f_mid, tol;
    int levels, ilo, ihi;
};

// Explicit-stack depth-first walk: two new abscissae per step; stop on levels <= 0 or
// |S2 - S| <= 15 tol (leaf value S2 + (S2-S)/15); else left half then right half with tol/2.
// `f(abscissa, position)`; positions are exact integers in units of width/2^26.  When
// `have_top`, the two new points of the first interval are supplied by the caller.
template <class F>
__device__ __forceinline__ double simpson_walk(F&& f, Span cur, bool have_top, double fd0, double fe0,
                                               Span* stack)
{
    double total = 0.0;
    int sp = 0;
    for (;;) {
        const double h = cur.hi - cur.lo;
        const double c = (cur.hi + cur.lo) / 2.;
        const int ic = (cur.ilo + cur.ihi) >> 1;
        double fd, fe;
        if (have_top) {
            fd = fd0;
            fe = fe0;
            have_top = false;
        } else {
            // one inlined copy of f serves both new abscissae (the walks run in the instruction-fetch-bound
            // part of the kernels: code size matters more than the loop overhead)
            fd = fe = 0.0;
#pragma unroll 1
            for (int side = 0; side < 2; ++side) {
                const double at = side ? (c + cur.hi) / 2. : (cur.lo + c) / 2.;
                const int pos = side ? (ic + cur.ihi) >> 1 : (cur.ilo + ic) >> 1;
                const double val = f(at, pos);
                if (side) fe = val; else fd = val;
            }
        }
        const double left = (h / 12.0) * ((cur.f_lo + 4.0 * fd) + cur.f_mid);
        const double right = (h / 12.0) * ((cur.f_mid + 4.0 * fe) + cur.f_hi);
        const double fine = left + right;
        if (cur.levels <= 0 || fabs(fine - cur.coarse) <= 15.0 * cur.tol) {
            total += fine + (fine - cur.coarse) / 15.0;
            if (sp == 0) break;
            cur = stack[--sp];
        } else {
            Span& r = stack[sp++];
            r.lo = c; r.hi = cur.hi; r.coarse = right; r.f_lo = cur.f_mid; r.f_hi = cur.f_hi; r.f_mid = fe;
            r.tol = cur.tol / 2; r.levels = cur.levels - 1; r.ilo = ic; r.ihi = cur.ihi;
            cur.hi = c; cur.coarse = left; cur.f_hi = cur.f_mid; cur.f_mid = fd;
            cur.tol = cur.tol / 2; cur.levels = cur.levels - 1; cur.ihi = ic;
        }
    }
    return total;
}

// Rare: the inner z-integration did not converge on its first level.
template <int KMAX>
__device__ __noinline__ double refine_z(const TrialCtx& c, const WTable<KMAX>& T, const TimeNode& n,
                                        const double f[5], double coarse, double tol, int levels,
                                        unsigned long long& evals)
{
    Span stack[kMaxDepth];
    Span top;
    top.lo = T.w[0]; top.hi = T.w[16]; top.coarse = coarse;
    top.f_lo = f[0]; top.f_hi = f[4]; top.f_mid = f[2];
    top.tol = tol; top.levels = levels; top.ilo = 0; top.ihi = kPosWidth;
    unsigned long long extra = 0;
    auto g = [&](double w, int pos) {
        extra++;
        const int m = (pos & ((1 << kPosShift) - 1)) == 0 ? (pos >> kPosShift) : -1;
        return node_value<KMAX>(c, T, n, w, m);
    };
    const double r = simpson_walk(g, top, true, f[1], f[3], stack);
    evals += extra;
    return r;
}

// adaptiveSimpsons_1D over z at decision time x_t (src/integrate.pxi:114-141), already / ZT.
template <int KMAX>
__device__ __noinline__ double inner_z(const TrialCtx& c, const WTable<KMAX>& T, double x_t, double tol,
                                       int n_sz, unsigned long long& evals)
{
    evals += 5;
    double f[5];
    if (x_t <= 0.0) {  // src/pdf.pxi:92-93: every node is exactly 0
        if (!(n_sz <= 0 || 0.0 <= 15.0 * tol))
            evals += (4ull << n_sz) - 4;  // tol < 0 or NaN: the reference walks the full tree of zeros
        return 0.0;
    }
    const TimeNode n = make_time_node(c, x_t);
    first_level<KMAX>(c, T, n, f);
    const double h = T.w[16] - T.w[0];
    const double coarse = (h / 6.0) * ((f[0] + 4.0 * f[2]) + f[4]);
    const double left = (h / 12.0) * ((f[0] + 4.0 * f[1]) + f[2]);
    const double right = (h / 12.0) * ((f[2] + 4.0 * f[3]) + f[4]);
    const double fine = left + right;
    if (n_sz <= 0 || fabs(fine - coarse) <= 15.0 * tol) return fine + (fine - coarse) / 15.0;
    return refine_z<KMAX>(c, T, n, f, coarse, tol, n_sz, evals);
}

// pdf_sv at the set's own start point (no z-integration): table node 8 is exactly z.
template <int KMAX>
__device__ __forceinline__ double single_z(const TrialCtx& c, const WTable<KMAX>& T, double x_t,
                                           unsigned long long& evals)
{
    evals++;
    if (x_t <= 0.0) return 0.0;
    const TimeNode n = make_time_node(c, x_t);
    return node_value<KMAX>(c, T, n, T.w[8], 8);
}

// full_pdf (src/pdf.pxi:104-146).  z and sz are the CTA's set-uniform values (they built WT);
// everything else may differ per trial.
template <int KMAX>
__device__ __forceinline__ double full_pdf_fast(double x, double v, double sv, double a, double z, double sz,
                                                double t, double st, const Numerics& nm,
                                                const SeriesTable* tab, const WTable<KMAX>* WT,
                                                unsigned long long& evals)
{
    if ((z < 0) || (z > 1) || (a < 0) || (t < 0) || (st < 0) || (sv < 0) || (sz < 0) || (sz > 1) ||
        ((fabs(x) - (t - st / 2.)) < 0) || (z + sz / 2. > 1) || (z - sz / 2. < 0) || (t - st / 2. < 0))
        return 0.0;
    const int b = x > 0 ? 1 : 0;
    if (b) v = -v;
    x = fabs(x);
    if (st < 1e-3) st = 0;
    if (sz < 1e-3) sz = 0;
    const WTable<KMAX>& T = WT[b];

    TrialCtx c;
    c.a_sq = a * a;
    c.inv_a_sq = 1.0 / c.a_sq;
    c.sv_sq = sv * sv;
    c.v_sq = v * v;
    c.asv = a * sv;
    c.av2 = (2.0 * a) * v;
    c.tab = tab;
    c.sv_zero = sv == 0.0;
    c.inv_zt = 1.0;

    if (sz == 0) {
        if (st == 0) return single_z<KMAX>(c, T, x - t, evals);  // :129
        const double lo = t - st / 2., hi = t + st / 2.;          // :131-134
        const double width = hi - lo;
        auto g = [&](double tn, int) { return single_z<KMAX>(c, T, x - tn, evals) / width; };
        Span top;
        top.lo = lo; top.hi = hi;
        top.f_lo = g(lo, 0); top.f_hi = g(hi, 0); top.f_mid = g((lo + hi) / 2., 0);
        top.coarse = (width / 6.0) * ((top.f_lo + 4.0 * top.f_mid) + top.f_hi);
        top.tol = nm.simps_err; top.levels = nm.n_st; top.ilo = 0; top.ihi = kPosWidth;
        Span stack[kMaxDepth];
        return simpson_walk(g, top, false, 0.0, 0.0, stack);
    }

    c.inv_zt = 1.0 / (T.w[16] - T.w[0]);
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        const double w = T.w[4 * i], t1 = c.asv * w;
        c.c5[i] = t1 * t1 - c.av2 * w;
    }
    if (st == 0) return inner_z<KMAX>(c, T, x - t, nm.simps_err, nm.n_sz, evals);  // :137-141

    const double lo = t - st / 2., hi = t + st / 2.;  // :143-146
    const double width = hi - lo;
    auto g = [&](double tn, int) { return inner_z<KMAX>(c, T, x - tn, nm.simps_err, nm.n_sz, evals) / width; };
    Span top;
    top.lo = lo; top.hi = hi;
    top.f_lo = g(lo, 0); top.f_hi = g(hi, 0); top.f_mid = g((lo + hi) / 2., 0);
    top.coarse = (width / 6.0) * ((top.f_lo + 4.0 * top.f_mid) + top.f_hi);
    top.tol = nm.simps_err; top.levels = nm.n_st; top.ilo = 0; top.ihi = kPosWidth;
    Span stack[kMaxDepth];
    return simpson_walk(g, top, false, 0.0, 0.0, stack);
}

}  // namespace hddm
