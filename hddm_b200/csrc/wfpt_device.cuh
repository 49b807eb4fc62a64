// wfpt_device.cuh -- FP64 device functions for the Wiener first-passage-time density.
//
// From-scratch CUDA (sm_100a) design of the math in the reference's src/pdf.pxi and
// src/integrate.pxi (file:line cited per function, relative to /root/reference).  Not a
// translation: the reference evaluates pdf_sv(x - t', v, sv, a, w, err) as a black box at every
// Simpson node; here everything that depends only on the decision time x_t = x - t' (normalised
// time, the Navarro-Fuss branch and term count K, the series scale, the sv-marginal
// denominators) is computed ONCE per time node in a NodeCtx and reused by every start-point
// node w of the inner z-integration, and the recursion is an explicit-stack depth-first walk.
// The DECISIONS (branch, K, Simpson stop tests, recursion tree) are exactly the reference's;
// only last-ulp libm / FMA rounding differs.
#pragma once
#include <cstdint>
#include <cuda_runtime.h>

#include "kernels.cuh"

namespace hddm {

constexpr double kPi = 3.14159265358979323846;
constexpr double kPiSq = 9.869604401089358;  // pow(M_PI, 2.0) in double

// Per-(trial, parameter set) constants after the boundary flip of src/pdf.pxi:116-120.
struct Core {
    double v, sv, a, a_sq, err;
    double sv_sq, v_sq;
    const SeriesTable* tab;    // shared-memory copy of the launch's table
    unsigned long long evals;  // pdf_sv evaluations made (the reference's unit of work)
};

// Navarro-Fuss series selection for one normalised time tt (src/pdf.pxi:36-54,60).
struct Series {
    double arg_scale;  // small-time: 0.5/tt        large-time: -pi^2 tt / 2
    double out_scale;  // small-time: 1/sqrt(2 pi tt^3)   large-time: pi
    int k_lo, k_hi;    // inclusive term range
    bool small_time;
};

// (branch, K) from the host-built table: +K small-time, <= 0 large-time with -code terms.
__device__ __forceinline__ int series_code(const SeriesTable& T, double tt)
{
    int code = T.code[0];
    for (int i = 0; i < T.n_thr; ++i)
        if (tt >= T.thr[i]) code = T.code[i + 1];
    return code;
}

// The reference's formulas evaluated on the device: only when the host could not build a table
// for this err (more than kMaxRegions regions, K out of range, non-finite err).
__device__ __noinline__ int series_code_direct(double tt, double err)
{
    double kl, ks;
    const double root_tt = sqrt(tt);
    const double kl_floor = 1. / (kPi * root_tt);
    const double pte = kPi * tt * err;
    if (pte < 1.0) {  // src/pdf.pxi:36-38
        kl = sqrt(-2.0 * log(pte) / (kPiSq * tt));
        kl = kl < kl_floor ? kl_floor : kl;
    } else {  // :39-40
        kl = kl_floor;
    }
    const double ste = 2.0 * sqrt(2.0 * kPi * tt) * err;
    if (ste < 1.0) {  // :43-45
        ks = 2.0 + sqrt(-2.0 * tt * log(ste));
        const double ks_floor = root_tt + 1.0;
        ks = ks < ks_floor ? ks_floor : ks;
    } else {  // :46-47
        ks = 2.0;
    }
    if (ks < kl) return (int)ceil(ks);  // :51-52
    return -(int)ceil(kl);              // :59-60
}

__device__ __forceinline__ Series select_series(double tt, const Core& c)
{
    Series s;
    const int code = c.tab->n_thr >= 0 ? series_code(*c.tab, tt) : series_code_direct(tt, c.err);
    if (code > 0) {  // src/pdf.pxi:51-54
        const int K = code;
        s.small_time = true;
        s.k_lo = -((K - 1) / 2);  // -floor((K-1)/2.)
        s.k_hi = K / 2;           // ceil((K-1)/2.)
        s.arg_scale = 0.5 / tt;
        s.out_scale = 1.0 / sqrt((2.0 * kPi) * (tt * tt * tt));
    } else {  // :59-60
        s.small_time = false;
        s.k_lo = 1;
        s.k_hi = -code;
        s.arg_scale = -kPiSq * tt / 2.0;
        s.out_scale = kPi;
    }
    return s;
}

// f(tt | 0, 1, w) given the selected series (src/pdf.pxi:55-57 / :61-63); ascending k.
__device__ __forceinline__ double series_density(const Series& s, double w)
{
    double p = 0.0;
    if (s.small_time) {
        for (int k = s.k_lo; k <= s.k_hi; ++k) {
            const double u = w + (double)(2 * k);
            p += u * exp(-(u * u) * s.arg_scale);
        }
    } else {
        for (int k = s.k_lo; k <= s.k_hi; ++k) {
            const double kd = (double)k;
            p += (kd * exp((kd * kd) * s.arg_scale)) * sin((kd * kPi) * w);
        }
    }
    return p * s.out_scale;
}

// Everything pdf_sv (src/pdf.pxi:87-102) needs that depends on x_t but not on w.
struct NodeCtx {
    Series ser;
    double inv_den;  // sv>0: 1/(2 sv^2 x_t + 2)
    double vx;       // v^2 x_t
    double out;      // sv>0: 1/(sqrt(sv^2 x_t + 1) a^2)   sv==0: 1/a^2
    bool live;       // x_t > 0 (src/pdf.pxi:92-93: otherwise the density is exactly 0)
};

__device__ __forceinline__ NodeCtx make_node(const Core& c, double x_t)
{
    NodeCtx n;
    n.live = !(x_t <= 0.0);  // NaN falls through as in the reference (x<=0 is false for NaN)
    if (n.live) {
        n.ser = select_series(x_t / c.a_sq, c);
        n.vx = c.v_sq * x_t;
        if (c.sv == 0.0) {
            n.inv_den = 0.0;
            n.out = 1.0 / c.a_sq;
        } else {
            n.inv_den = 1.0 / (2.0 * c.sv_sq * x_t + 2.0);
            n.out = 1.0 / (sqrt(c.sv_sq * x_t + 1.0) * c.a_sq);
        }
    }
    return n;
}

// pdf_sv at start point w for a prepared time node.  src/pdf.pxi:85 (sv==0) and :102 (sv>0,
// exp(log(p)+E) == p*exp(E) for p>0; p<0 gives NaN, p==0 gives 0 unless E is NaN).
__device__ __forceinline__ double node_density(Core& c, const NodeCtx& n, double w)
{
    c.evals++;
    if (!n.live) return 0.0;
    const double p = series_density(n.ser, w);
    if (c.sv == 0.0) return p * exp(-c.v * c.a * w - n.vx / 2.) * n.out;
    if (p < 0.0) return __longlong_as_double(0x7ff8000000000000LL);
    const double azs = c.a * w * c.sv;
    const double e = (azs * azs - 2.0 * c.a * c.v * w - n.vx) * n.inv_den;
    return p * exp(e) * n.out;
}

__device__ __forceinline__ double pdf_sv_at(Core& c, double x_t, double w)
{
    const NodeCtx n = make_node(c, x_t);
    return node_density(c, n, w);
}

// src/pdf.pxi:67-72
__device__ __forceinline__ double prob_ub(double v, double a, double z)
{
    if (v == 0.0) return z;
    return (exp(-2.0 * a * z * v) - 1.0) / (exp(-2.0 * a * v) - 1.0);
}

// ---------------------------------------------------------------------------------------------
// Adaptive Simpson (src/integrate.pxi:72-141), explicit stack.  `F` maps an abscissa to the
// already width-normalised integrand.  The walk is the reference's: two new points per step,
// stop on levels<=0 or |S2-S| <= 15*tol, else left half then right half with tol/2; leaves
// return S2 + (S2-S)/15.  Leaf values are accumulated left to right.
struct Pending {
    double lo, hi, coarse, f_lo, f_hi, f_mid, tol;
    int levels;
};

template <class F>
__device__ __forceinline__ double adaptive_simpson(F&& f, double lo, double hi, double tol, int depth,
                                                   Pending* stack)
{
    const double h0 = hi - lo;
    double f_lo = f(lo);
    double f_hi = f(hi);
    double f_mid = f((lo + hi) / 2.);
    double coarse = (h0 / 6.0) * ((f_lo + 4.0 * f_mid) + f_hi);
    double total = 0.0;
    int sp = 0, levels = depth;
    for (;;) {
        const double h = hi - lo;
        const double c = (hi + lo) / 2.;
        const double fd = f((lo + c) / 2.);
        const double fe = f((c + hi) / 2.);
        const double left = (h / 12.0) * ((f_lo + 4.0 * fd) + f_mid);
        const double right = (h / 12.0) * ((f_mid + 4.0 * fe) + f_hi);
        const double fine = left + right;
        if (levels <= 0 || fabs(fine - coarse) <= 15.0 * tol) {
            total += fine + (fine - coarse) / 15.0;
            if (sp == 0) break;
            const Pending& q = stack[--sp];
            lo = q.lo; hi = q.hi; coarse = q.coarse; f_lo = q.f_lo; f_hi = q.f_hi; f_mid = q.f_mid;
            tol = q.tol; levels = q.levels;
        } else {
            Pending& q = stack[sp++];
            q.lo = c; q.hi = hi; q.coarse = right; q.f_lo = f_mid; q.f_hi = f_hi; q.f_mid = fe;
            q.tol = tol / 2; q.levels = levels - 1;
            hi = c; coarse = left; f_hi = f_mid; f_mid = fd;
            tol = tol / 2; levels = levels - 1;
        }
    }
    return total;
}

// Fixed composite Simpson with the reference's quirks (src/integrate.pxi:12-45): weights
// 1,4,2,4,...,2 then subtract the last node once; divide by the interval; times h/3.
template <class F>
__device__ __forceinline__ double fixed_simpson(F&& f, double lo, double hi, int n)
{
    const double h = (hi - lo) / n;
    double acc = f(lo), y = 0.0;
    for (int i = 1; i <= n; ++i) {
        y = f(lo + h * i);
        acc = (i & 1) ? acc + 4.0 * y : acc + 2.0 * y;
    }
    acc = acc - y;
    acc = acc / (hi - lo);
    return (h * acc) / 3.0;
}

// full_pdf (src/pdf.pxi:104-146): support test on the UNFLIPPED parameters, boundary flip,
// st/sz < 1e-3 -> 0, then the four-way dispatch.  `evals` returns the number of pdf_sv calls.
static __device__ __noinline__ double full_pdf(double x, double v, double sv, double a, double z, double sz,
                                        double t, double st, const Numerics& nm,
                                        const SeriesTable* tab, unsigned long long& evals)
{
    if ((z < 0) || (z > 1) || (a < 0) || (t < 0) || (st < 0) || (sv < 0) || (sz < 0) || (sz > 1) ||
        ((fabs(x) - (t - st / 2.)) < 0) || (z + sz / 2. > 1) || (z - sz / 2. < 0) || (t - st / 2. < 0))
        return 0.0;
    if (x > 0) {
        v = -v;
        z = 1. - z;
    }
    x = fabs(x);
    if (st < 1e-3) st = 0;
    if (sz < 1e-3) sz = 0;

    Core c;
    c.v = v; c.sv = sv; c.a = a; c.a_sq = a * a; c.err = nm.err;
    c.sv_sq = sv * sv; c.v_sq = v * v; c.tab = tab; c.evals = 0;
    double out;

    if (sz == 0 && st == 0) {  // :129
        out = pdf_sv_at(c, x - t, z);
    } else if (sz == 0) {  // :131-134, integrate over t at fixed z
        const double lo = t - st / 2., hi = t + st / 2.;
        const double width = hi - lo;
        auto f = [&](double tn) { return pdf_sv_at(c, x - tn, z) / width; };
        if (nm.use_adaptive > 0) {
            Pending stack[kMaxDepth];
            out = adaptive_simpson(f, lo, hi, nm.simps_err, nm.n_st, stack);
        } else {
            auto g = [&](double tn) { return pdf_sv_at(c, x - tn, z); };
            out = fixed_simpson(g, lo, hi, nm.n_st);
        }
    } else if (st == 0) {  // :137-141, integrate over z at fixed t: ONE time node for all w
        const double lo = z - sz / 2., hi = z + sz / 2.;
        const double width = hi - lo;
        const NodeCtx node = make_node(c, x - t);
        if (nm.use_adaptive) {
            auto f = [&](double w) { return node_density(c, node, w) / width; };
            Pending stack[kMaxDepth];
            out = adaptive_simpson(f, lo, hi, nm.simps_err, nm.n_sz, stack);
        } else {
            auto g = [&](double w) { return node_density(c, node, w); };
            out = fixed_simpson(g, lo, hi, nm.n_sz);
        }
    } else {  // :143-146, outer over t, inner over z
        const double zlo = z - sz / 2., zhi = z + sz / 2.;
        const double tlo = t - st / 2., thi = t + st / 2.;
        const double zwidth = zhi - zlo;
        const double twidth = thi - tlo;
        if (nm.use_adaptive) {
            Pending zstack[kMaxDepth];
            Pending tstack[kMaxDepth];
            auto inner = [&](double tn) {
                const NodeCtx node = make_node(c, x - tn);
                auto f = [&](double w) { return node_density(c, node, w) / zwidth; };
                return adaptive_simpson(f, zlo, zhi, nm.simps_err, nm.n_sz, zstack) / twidth;
            };
            out = adaptive_simpson(inner, tlo, thi, nm.simps_err, nm.n_st, tstack);
        } else {
            auto inner = [&](double tn) {
                const NodeCtx node = make_node(c, x - tn);
                auto g = [&](double w) { return node_density(c, node, w); };
                return fixed_simpson(g, zlo, zhi, nm.n_sz);
            };
            out = fixed_simpson(inner, tlo, thi, nm.n_st);
        }
    }
    evals += c.evals;
    return out;
}

}  // namespace hddm
