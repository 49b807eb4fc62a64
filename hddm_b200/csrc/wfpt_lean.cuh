// wfpt_lean.cuh -- the first 5 x 5 level of full_pdf (src/pdf.pxi:104-146 with src/integrate.pxi:143-206) for the
// headline case, written for the FP64 pipe: a warp whose 32 trials all sit in the large-time regime of the
// Navarro-Fuss series (src/pdf.pxi:59-63) at all five decision times, full DDM (sv, sz, st on), drift uniform per
// parameter set.  Everything else -- small-time nodes, a node outside the comfortable range, a Simpson stop
// test that fails, a negative series value -- is left to the evaluators of wfpt_grid.cuh (same warp round, or
// deferred to the second pass), so this file never changes a DECISION of the reference; it only spends fewer
// instructions on the ~95 % of the headline sweep's trials that need nothing but the first level.
//
// What is different from grid_node / full_pdf_grid<.., 1>  (profiles/r2_v1_*: 521 FP64 + 610 other warp
// instructions per trial, FP64 pipe 57.7 % busy, issue-slot bound):
//   * the number of series terms K is a WARP-uniform template parameter (trials are laid out by |rt|, so a
//     warp's nodes share K except at the few break points; there the extra term of a node with fewer terms gets
//     a zero coefficient by a predicated select) -- no per-node branching, no (branch, K) search per node;
//   * large-time term k carries q^(k^2) = q * q^(k^2 - 1): the common q rides on the scale of the sv-marginal
//     exponential and k sin(k pi w_i) comes from the per-set table, so a K-term node costs 5 (K - 1) FMAs;
//   * pi / a^2, both Simpson h/12 factors and both pre-divisions by the interval widths (src/integrate.pxi:129,
//     134, 187-198) are one per-set constant k1 = pi / (144 a^2): with phi'' = phi / 144 the inner stop test reads
//     |D phi''| <= 15 simps_err / 12 and the outer one |D I''| <= 15 simps_err, where for five equally spaced
//     values D = 4 (y1 + y3) - (y0 + y4) - 6 y2 is 12 (fine - coarse) / h and S = (y0 + y4) + 4 (y1 + y3) + 2 y2 is
//     12 fine / h; the accepted value is S + D / 15 (src/integrate.pxi:99-106);
//   * exp through a 2048-entry table of 2^(i/2048) in shared memory and a cubic (7 FP64 instructions, 4 integer
//     ones; relative error < 1e-15 for |x| <= 128, which the range checks below guarantee).
#pragma once
#include "wfpt_grid.cuh"

namespace hddm {

#ifndef HDDM_LEAN_UNROLL
#define HDDM_LEAN_UNROLL 5
#endif
constexpr int kLeanUnroll = HDDM_LEAN_UNROLL;  // the five decision-time nodes of a lean trial: unrolled (5) or looped (1)
constexpr int kExpBigBits = 11;
constexpr int kExpBig = 1 << kExpBigBits;
// 2^(i/2048), filled once per process by the host (launch_like -> ensure_exp_table) with correctly rounded values.
__device__ double g_exp2_big[kExpBig];
#ifdef HDDM_DEBUG_TREE
__device__ unsigned long long g_dbg[16];
#define DBG(i, c) do { if (c) atomicAdd(&g_dbg[i], 1ULL); } while (0)
#else
#define DBG(i, c) do { } while (0)
#endif

// exp(x), |x| <= 128 (callers guarantee it): x = n ln2/2048 + r, |r| <= ln2/4096 = 1.7e-4;
// exp(x) = 2^(n >> 11) T[n & 2047] (1 + r + r^2/2 + r^3/6), truncation r^4/24 < 4e-17, argument reduction
// with the single constant ln2/2048 (its rounding error times |n| <= 3.8e5 stays below 1.5e-14 |x| / 128).
// CLAMP: any x <= 128 down to -7e5; results below 2^-1021 come out as exactly 0 (used where such a term is added to
// one larger by hundreds of orders of magnitude, or multiplies a node that is negligible as a whole).
template <bool CLAMP = false>
__device__ __forceinline__ double exp_lean(double x, const double* __restrict__ T)
{
    const double magic = 6755399441055744.0;  // 1.5 * 2^52
    const double t = fma(x, 0x1.71547652b82fep+11, magic);
    const int n = __double2loint(t);
    const double nd = t - magic;
    const double r = fma(nd, -0x1.62e42fefa39efp-12, x);
    const double c = fma(r, 1.0 / 6.0, 0.5);
    const double p = fma(r * r, c, r);
    const double tv = T[n & (kExpBig - 1)];
    const double m = fma(tv, p, tv);
    // 2^(n >> 11) goes into the exponent field: (n >> 11) << 20 == (n with its 11 index bits cleared) * 2^9, one
    // mask and one multiply-add
    int hi = (n & -kExpBig) * (1 << (20 - kExpBigBits)) + __double2hiint(m), lo = __double2loint(m);
    if (CLAMP && n < -1021 * kExpBig) hi = lo = 0;
    return __hiloint2double(hi, lo);
}

// Fills L from the GridSet it belongs to; one thread (tid == 0) after build_gridset's final barrier.
template <class GS>
__device__ __noinline__ void build_leanset(GS& G, LeanSet& L, double v, const Numerics& nm, bool drift_uniform)
{
    constexpr unsigned long long kInf = 0x7ff0000000000000ULL;
    for (int b = 0; b < 2; ++b) {
        for (int k = 0; k < 4; ++k)
            for (int i = 0; i < 6; ++i) L.ks[b][k][i] = (double)(k + 1) * G.s5[b][k][i];
        const double vb = b ? -v : v;  // src/pdf.pxi:116-118
        L.c0h[b] = 0.5 * fma(-G.aw0[b], vb, G.sw0[b]);
        L.dc0h[b] = 0.5 * fma(-G.daw[b], vb, G.dsw[b]);
        L.d2ch[b] = 0.5 * G.d2c[b];
    }
    L.v2h = 0.5 * (v * v);
    L.v = v;
    L.k1 = kPi / (144.0 * G.a_sq);
    L.thr_out = 15.0 * nm.simps_err;
    L.thr_in = L.thr_out / 12.0;
    // Break points: regions of the window ascend in decision time; the lean level needs the tail of the window to
    // be large-time with a non-increasing term count of at most 4 (true for every err the table builder accepts
    // with max_large_k <= 4; checked, not assumed).
    int n_win = 0;
    while (n_win < kThrWin - 1 && G.xwin[n_win] != kInf) ++n_win;   // thresholds inside the window
    int s = 0;
    while (s <= n_win && G.code[G.region0 + s] > 0) ++s;             // first large-time region
    bool mono = s <= n_win;
    for (int m = s; m <= n_win && mono; ++m) {
        const int c = G.code[G.region0 + m];
        mono = c < 0 && -c <= 4 && (m == s || -c <= -G.code[G.region0 + m - 1]);
    }
    const unsigned long long edge = s == 0 ? 0ULL : G.xwin[s - 1];
    for (int k = 2; k <= 4; ++k) {
        // first region (from s) with fewer than k terms: xt >= its lower edge <=> fewer than k terms
        int m = s;
        while (m <= n_win && -G.code[G.region0 + m] >= k) ++m;
        L.x_k[k - 2] = m == s ? edge : (m > n_win ? kInf : G.xwin[m - 1]);
    }
    for (int b = 0; b < 2; ++b) {
        const unsigned long long lo = (unsigned long long)__double_as_longlong(G.xt_lo[b]);
        L.x_small[b] = (long long)(edge > lo ? edge : lo);
    }
    // exponent arguments: -pi^2 tt / 2 >= -128 and (v^2 xt + 2 a |v|) / 2 + (a sv w_4)^2 / 2 <= 128 (GridSet::ok bounds
    // the last term by 300: tightened here)
    const double wmax = G.wt[0].w[16] > G.wt[1].w[16] ? G.wt[0].w[16] : G.wt[1].w[16];
    const double asw = (G.a * G.sv) * wmax;
    double xhi = 25.0 * G.a_sq;
    const double room = 200.0 - (2.0 * G.a) * fabs(v);
    if (drift_uniform && v != 0.0 && room / (v * v) < xhi) xhi = room / (v * v);  // (per-trial drift: lean_begin_v)
    if (G.xt_hi < xhi) xhi = G.xt_hi;
    L.xhi = xhi;
    // small-time side of the window: 3 terms, then 4 terms (err = 1e-4: 3 below tt = 0.0555, 4 up to 0.154)
    bool small_ok = s >= 1;
    int first4 = s;  // first small-time region with 4 terms
    for (int m = 0; m < s && small_ok; ++m) {
        const int c = G.code[G.region0 + m];
        small_ok = (c == 3 && first4 == s) || c == 4;
        if (c == 4 && first4 == s) first4 = m;
    }
    L.x_edge = edge;
    L.x_s4 = first4 == s ? edge : (first4 == 0 ? 0ULL : G.xwin[first4 - 1]);
    for (int b = 0; b < 2; ++b) {
        // down to the comfortable range's lower end (the k = 0 term stays above e^-650 there; the others may
        // underflow: exp_lean<true>)
        L.x_lo_small[b] = __double_as_longlong(G.xt_lo[b]);
        L.xz[b] = nm.simps_err >= 1e-200 ? G.xt_zero[b] : 0.0;
        small_ok = small_ok && G.chain_ok[b] && (unsigned long long)L.x_lo_small[b] < edge;
        L.g_w0sq[b] = (0.5 * G.a_sq) * G.w0sq[b];
        L.g_dw1[b] = (0.5 * G.a_sq) * G.dw1[b];
        L.g_d2w[b] = (0.5 * G.a_sq) * G.d2w[b];
        L.g4_omw4[b] = (2.0 * G.a_sq) * G.omw4[b];
        L.g4_w0p1[b] = (2.0 * G.a_sq) * G.w0p1[b];
    }
    L.g4_hw = (2.0 * G.a_sq) * G.hw;
    L.g4_one = 2.0 * G.a_sq;
    L.k1s = 1.0 / ((144.0 * G.a_sq) * sqrt(2.0 * kPi));
    L.uniform = drift_uniform ? 1 : 0;
    const bool inner_ok = G.ok && G.support && mono && G.sz != 0 && nm.n_sz > 0 && nm.simps_err >= 0.0 &&
                          asw * asw <= 50.0 && xhi > 0.0 && (!drift_uniform || isfinite(v)) && isfinite(L.k1) && L.k1 > 0.0;
    L.ok = inner_ok && G.st != 0 && nm.n_st > 0 ? 1 : 0;
    L.ok_small = L.ok && small_ok && isfinite(L.k1s) ? 1 : 0;
    L.ok1 = inner_ok && G.st == 0 && drift_uniform ? (small_ok && isfinite(L.k1s) ? 3 : 1) : 0;
    // neither sz nor st (after the snaps of src/pdf.pxi:122-125): full_pdf is pdf_sv at the one start point and the one
    // decision time (src/pdf.pxi:127-128) -- lean_simple; bits 2 (applies) and 3 (to small-time decision times too)
    const bool point_ok = G.ok && G.support && mono && G.sz == 0 && G.st == 0 && asw * asw <= 50.0 && xhi > 0.0 &&
                          drift_uniform && isfinite(v) && isfinite(L.k1) && L.k1 > 0.0;
    if (point_ok) L.ok1 |= small_ok && isfinite(L.k1s) ? 12 : 4;
    L.tn17[0] = G.tn[0];
    L.tn17[16] = G.tn[4];
    for (int step = 8; step >= 1; step >>= 1)  // (lb + ub) / 2. recursively: tn17[4 j] == tn[j] bit for bit
        for (int m = step; m < 16; m += 2 * step) L.tn17[m] = (L.tn17[m - step] + L.tn17[m + step]) / 2.;
    L.pad17 = 0.0;
}

__device__ __forceinline__ bool below(double xt, unsigned long long edge)
{
    return (unsigned long long)__double_as_longlong(xt) < edge;
}

// The same for a decision time known to be positive (or a lane whose result is discarded), as one FP64 compare
// instead of two integer ones: positive doubles order like their bit patterns, +inf as an edge takes every finite xt,
// NaN fails both.
__device__ __forceinline__ bool below_pos(double xt, unsigned long long edge)
{
    return xt < __longlong_as_double((long long)edge);
}

// Per-trial constants of the lean evaluators (boundary-dependent table entries, read once per trial).
struct LeanTrial {
    int b;
    double c0h, dc0h, d2ch;
    double v2h, vb;  // v^2 / 2 and the drift after the boundary flip (src/pdf.pxi:116-118)
};

// sv marginal of one decision time (src/pdf.pxi:102) times `scale`, at the five start points:
// exp(c(w_i) d - v^2 xt d), d = r^2 / 2 -- the numerator is quadratic in i, so the five exponentials are a
// second-order geometric chain from three.  `shift0/1/2` are added to the three exponents (the small-time node
// merges its own quadratic chain exp(-w_i^2 / (2 tt)) here).
template <bool CLAMP>
__device__ __forceinline__ void lean_chain(const LeanSet& L, const LeanTrial& c, const double* __restrict__ T, double xt,
                                           double dd, double scale, double shift0, double shift1, double shift2,
                                           double E[5])
{
    E[0] = exp_lean<CLAMP>(fma(fma(-c.v2h, xt, c.c0h), dd, shift0), T) * scale;
    const double g0 = exp_lean<CLAMP>(fma(c.dc0h, dd, shift1), T);
    const double r2 = exp_lean<CLAMP>(fma(c.d2ch, dd, shift2), T);
    E[1] = E[0] * g0;
    const double g1 = g0 * r2;
    E[2] = E[1] * g1;
    const double g2 = g1 * r2;
    E[3] = E[2] * g2;
    E[4] = E[3] * (g2 * r2);
}

// Inner Simpson level over the five start points (src/integrate.pxi:114-141, :99-106) of values p_i E_i.
__device__ __forceinline__ double lean_inner(const LeanSet& L, const double p[5], const double E[5], bool& bad)
{
    const double s04 = fma(p[4], E[4], p[0] * E[0]);
    const double s13 = fma(p[3], E[3], p[1] * E[1]);
    const double f2 = p[2] * E[2];
    const double S = fma(2.0, f2, fma(4.0, s13, s04));
    const double D = fma(-6.0, f2, fma(4.0, s13, -s04));
    bad |= !(fabs(D) <= L.thr_in);
    return fma(D, 1.0 / 15.0, S);
}

// The same, reporting the stop test instead of folding it into `bad`, and handing out the five values.
__device__ __forceinline__ double lean_inner_f(const LeanSet& L, const double p[5], const double E[5], bool* pass,
                                               double f[5])
{
#pragma unroll
    for (int i = 0; i < 5; ++i) f[i] = p[i] * E[i];
    const double s04 = f[0] + f[4], s13 = f[1] + f[3];
    const double S = fma(2.0, f[2], fma(4.0, s13, s04));
    const double D = fma(-6.0, f[2], fma(4.0, s13, -s04));
    *pass = fabs(D) <= L.thr_in;
    return fma(D, 1.0 / 15.0, S);
}

// One LARGE-TIME decision time xt (src/pdf.pxi:59-63) with q = exp(-pi^2 tt / 2): the inner integral I''.
// wK (warp-uniform): the largest term count among the warp's nodes; a node with fewer terms gets zero coefficients.
template <int KMAX>
__device__ __forceinline__ double lean_node_large(const GridSet<KMAX>& G, const LeanSet& L, const LeanTrial& c,
                                                  const double* __restrict__ T, double xt, double q, int wK, int& neg,
                                                  bool& bad)
{
    // series / q: sin(pi w_i) + sum_{k>=2} q^(k^2-1) k sin(k pi w_i), ascending k (src/pdf.pxi:60-62)
    // (the k = 2 term is evaluated whatever wK says: nine nodes in ten of the headline sweep have it, a node without
    //  it gets a zero coefficient, and the first sines, which the five nodes of a trial share, need no copy)
    double p[5], s[5];
    {
        double s1[5];
        load5(L.ks[c.b][0], s1);
        const double q2 = q * q, q3 = q2 * q;
        const double e2 = below_pos(xt, L.x_k[0]) ? q3 : 0.0;
        load5(L.ks[c.b][1], s);
#pragma unroll
        for (int i = 0; i < 5; ++i) p[i] = fma(e2, s[i], s1[i]);
        if (wK >= 3) {
            const double q4 = q2 * q2, q8 = q4 * q4;
            const double e3 = below_pos(xt, L.x_k[1]) ? q8 : 0.0;
            load5(L.ks[c.b][2], s);
#pragma unroll
            for (int i = 0; i < 5; ++i) p[i] = fma(e3, s[i], p[i]);
            if (wK >= 4) {
                const double e4 = below_pos(xt, L.x_k[2]) ? (q8 * q4) * q3 : 0.0;
                load5(L.ks[c.b][3], s);
#pragma unroll
                for (int i = 0; i < 5; ++i) p[i] = fma(e4, s[i], p[i]);
            }
        }
    }
    neg |= (__double2hiint(p[0]) | __double2hiint(p[1])) | (__double2hiint(p[2]) | __double2hiint(p[3])) |
           __double2hiint(p[4]);
    const double r = rsqrt_fast(fma(G.sv_sq, xt, 1.0));
    double E[5];
    lean_chain<false>(L, c, T, xt, r * r, (r * q) * L.k1, 0.0, 0.0, 0.0, E);
    return lean_inner(L, p, E, bad);
}

// One SMALL-TIME decision time xt (src/pdf.pxi:51-57) with 3 terms (k = -1, 0, 1) or, where `four`, 4 (k = 2 too).
// p_i = sum_k (w_i + 2k) exp(-(w_i + 2k)^2 g), g = 1 / (2 tt) = a^2 / (2 xt): exp(-(w_i + 2k)^2 g) = A_i R_{k,i} with
// A_i = exp(-g w_i^2) (quadratic in i: merged into the sv-marginal chain) and R_{k,i} = exp(-4 g k (w_i + k)),
// geometric in i with ratio eta = exp(-4 g h_w); R_{2,i} = R_{1,i}^2 exp(-8 g).
template <int KMAX>
__device__ __forceinline__ double lean_node_small(const GridSet<KMAX>& G, const LeanSet& L, const LeanTrial& c,
                                                  const double* __restrict__ T, double xt, bool any_four, int& neg,
                                                  bool& bad, bool* inner_pass = nullptr, double* f5 = nullptr)
{
    const int b = c.b;
    const double rc = rcp_fast(xt);
    double w[5];
    load5(G.w5[b], w);
    const double g4 = L.g4_one * rc;
    const double eta = exp_lean<true>(-L.g4_hw * rc, T);
    double Rn = exp_lean<true>(-L.g4_omw4[b] * rc, T);   // k = -1 at i = 4, then down
    double Rp = exp_lean<true>(-L.g4_w0p1[b] * rc, T);   // k = +1 at i = 0, then up
    double C2 = 0.0;
    if (any_four) {  // warp-uniform
        const double C = exp_lean<true>(-g4, T);
        C2 = below(xt, L.x_s4) ? 0.0 : C * C;
    }
    double t[5];
#pragma unroll
    for (int i = 4; i >= 0; --i) {  // the reference adds k ascending: k = -1 first
        t[i] = fma(w[i] - 2.0, Rn, w[i]);
        if (i > 0) Rn *= eta;
    }
#pragma unroll
    for (int i = 0; i < 5; ++i) {
        t[i] = fma(w[i] + 2.0, Rp, t[i]);
        if (any_four) t[i] = fma(w[i] + 4.0, (Rp * Rp) * C2, t[i]);
        if (i < 4) Rp *= eta;
    }
    neg |= (__double2hiint(t[0]) | __double2hiint(t[1])) | (__double2hiint(t[2]) | __double2hiint(t[3])) |
           __double2hiint(t[4]);
    const double r = rsqrt_fast(fma(G.sv_sq, xt, 1.0));
    // 1 / sqrt(2 pi tt^3) = u sqrt(u) / sqrt(2 pi), u = 1 / tt = a^2 / xt
    const double u = G.a_sq * rc;
    const double post = (u * u) * rsqrt_fast(u);
    double E[5];
    lean_chain<true>(L, c, T, xt, r * r, (post * r) * L.k1s, -L.g_w0sq[b] * rc, -L.g_dw1[b] * rc, -L.g_d2w[b] * rc, E);
    if (f5) return lean_inner_f(L, t, E, inner_pass, f5);
    return lean_inner(L, t, E, bad);
}

// pdf_sv'' at ONE start point w of a small-time decision time (the deeper levels of the inner recursion,
// src/integrate.pxi:99-112): the same terms as lean_node_small without the chains.
template <int KMAX>
__device__ __forceinline__ double lean_point_small(const GridSet<KMAX>& G, const LeanSet& L, const LeanTrial& c,
                                                   const double* __restrict__ T, double xt, double w, double rc,
                                                   double dd, double scale, double C2, int& neg)
{
    const double g4 = L.g4_one * rc, g = 0.25 * g4;
    const double Rn = exp_lean<true>(-g4 * (1.0 - w), T), Rp = exp_lean<true>(-g4 * (w + 1.0), T);
    double t = fma(w - 2.0, Rn, w);
    t = fma(w + 2.0, Rp, t);
    t = fma(w + 4.0, (Rp * Rp) * C2, t);
    neg |= __double2hiint(t);
    const double asw = (G.a * G.sv) * w;
    const double ch = 0.5 * fma(asw, asw, -((2.0 * G.a) * c.vb) * w);
    return (t * scale) * exp_lean<true>(fma(fma(-c.v2h, xt, ch), dd, -(g * w) * w), T);
}

// The inner recursion below its first level for a small-time decision time whose stop test failed: f5 = the five
// first-level values (table nodes 0, 4, 8, 12, 16).  Same tree as the outer one (n_sz == 2).  *extra += pdf_sv made.
template <int KMAX>
__device__ __noinline__ double lean_refine_small(const GridSet<KMAX>& G, const LeanSet& L, const LeanTrial& c,
                                                 const double* __restrict__ T, double xt, const double f5[5], int& neg,
                                                 int* extra)
{
    const double rc = rcp_fast(xt);
    const double r = rsqrt_fast(fma(G.sv_sq, xt, 1.0));
    const double u = G.a_sq * rc;
    const double scale = (((u * u) * rsqrt_fast(u)) * r) * L.k1s, dd = r * r;
    double C2 = 0.0;
    if (!below(xt, L.x_s4)) {
        const double C = exp_lean<true>(-L.g4_one * rc, T);
        C2 = C * C;
    }
    const double* wn = G.wt[c.b].w;
    auto pt = [&](int m) { return lean_point_small<KMAX>(G, L, c, T, xt, wn[m], rc, dd, scale, C2, neg); };
    auto simpson = [&](double y0, double y1, double y2, double y3, double y4, bool* pass) {
        const double s04 = y0 + y4, s13 = y1 + y3;
        const double S = fma(2.0, y2, fma(4.0, s13, s04));
        const double D = fma(-6.0, y2, fma(4.0, s13, -s04));
        *pass = fabs(D) <= L.thr_in;
        return fma(D, 1.0 / 15.0, S);
    };
    double Z[17];
    Z[0] = f5[0]; Z[4] = f5[1]; Z[8] = f5[2]; Z[12] = f5[3]; Z[16] = f5[4];
#pragma unroll 1
    for (int m = 2; m < 16; m += 4) Z[m] = pt(m);
    int made = 4;
    bool lok, rok, dummy;
    double left = 0.5 * simpson(Z[0], Z[2], Z[4], Z[6], Z[8], &lok);
    double right = 0.5 * simpson(Z[8], Z[10], Z[12], Z[14], Z[16], &rok);
#pragma unroll 1
    for (int m = 1; m < 16; m += 2) {
        if (m < 8 ? lok : rok) continue;
        Z[m] = pt(m);
        ++made;
    }
    if (!lok) left = 0.25 * simpson(Z[0], Z[1], Z[2], Z[3], Z[4], &dummy) + 0.25 * simpson(Z[4], Z[5], Z[6], Z[7], Z[8], &dummy);
    if (!rok) right = 0.25 * simpson(Z[8], Z[9], Z[10], Z[11], Z[12], &dummy) + 0.25 * simpson(Z[12], Z[13], Z[14], Z[15], Z[16], &dummy);
    *extra += made;
    return left + right;
}

// Outer Simpson level over the five decision times (src/integrate.pxi:143-178).
__device__ __forceinline__ double lean_outer(const LeanSet& L, const double I[5], bool& bad)
{
    const double s04 = I[0] + I[4], s13 = I[1] + I[3];
    const double S = fma(2.0, I[2], fma(4.0, s13, s04));
    const double D = fma(-6.0, I[2], fma(4.0, s13, -s04));
    bad |= !(fabs(D) <= L.thr_out);
    return fma(D, 1.0 / 15.0, S);
}

__device__ __forceinline__ LeanTrial lean_begin(const LeanSet& L, int b)
{
    LeanTrial c;
    c.b = b;
    c.c0h = L.c0h[b];
    c.dc0h = L.dc0h[b];
    c.d2ch = L.d2ch[b];
    c.v2h = L.v2h;
    c.vb = b ? -L.v : L.v;
    return c;
}

// The same for a drift that differs from trial to trial (regression on v): the two drift-dependent chain constants
// are two FMAs per trial.  *fits: the exponent bound the per-set range check makes for a uniform drift.
template <int KMAX>
__device__ __forceinline__ LeanTrial lean_begin_v(const GridSet<KMAX>& G, const LeanSet& L, int b, double v, double xt0,
                                                  bool* fits)
{
    LeanTrial c;
    c.b = b;
    c.vb = b ? -v : v;
    c.c0h = 0.5 * fma(-G.aw0[b], c.vb, G.sw0[b]);
    c.dc0h = 0.5 * fma(-G.daw[b], c.vb, G.dsw[b]);
    c.d2ch = L.d2ch[b];
    c.v2h = 0.5 * (v * v);
    *fits = fma(v * v, xt0, (2.0 * G.a) * fabs(v)) <= 200.0;  // (NaN drift: false)
    return c;
}

// One trial, all five decision times large-time with at most wK terms (warp-uniform).  b: boundary (0 lower /
// 1 upper); ax = |rt|; xt0 = ax - tn[0] and xt4 = ax - tn[4] have passed the range checks.  Returns full_pdf; *need
// is set when a stop test fails or a series value is negative (the trial is then re-evaluated by the complete
// evaluator and the return value ignored).
template <int KMAX>
__device__ __forceinline__ double lean_trial_large(const GridSet<KMAX>& G, const LeanSet& L, const double* __restrict__ T,
                                                   const LeanTrial& c, double ax, double xt0, double xt4, int wK,
                                                   bool* need)
{
    const double q0 = exp_lean(G.qk * xt0, T);
    int neg = 0;
    bool bad = false;
    double out;
    if (kLeanUnroll == 5) {
        double I[5];
#pragma unroll
        for (int j = 0; j < 5; ++j) {
            const double xt = j == 0 ? xt0 : (j == 4 ? xt4 : ax - G.tn[j]);
            const double q = j == 0 ? q0 : q0 * G.rho[j];
            I[j] = lean_node_large<KMAX>(G, L, c, T, xt, q, wK, neg, bad);
        }
        out = lean_outer(L, I, bad);
    } else {  // looped: S'' and D'' of the outer level accumulate with the Simpson weights 1 4 2 4 1 / -1 4 -6 4 -1
        double S = 0.0, D = 0.0;
#pragma unroll 1
        for (int j = 0; j < 5; ++j) {
            const double xt = ax - G.tn[j];
            const double v = lean_node_large<KMAX>(G, L, c, T, xt, q0 * G.rho[j], wK, neg, bad);
            const double ws = (j & 1) ? 4.0 : (j == 2 ? 2.0 : 1.0), wd = (j & 1) ? 4.0 : (j == 2 ? -6.0 : -1.0);
            S = fma(ws, v, S);
            D = fma(wd, v, D);
        }
        bad |= !(fabs(D) <= L.thr_out);
        out = fma(D, 1.0 / 15.0, S);
    }
    *need = bad | (neg < 0);
    return out;
}

// The same with all five decision times small-time.
template <int KMAX>
__device__ __forceinline__ double lean_trial_small(const GridSet<KMAX>& G, const LeanSet& L, const double* __restrict__ T,
                                                   const LeanTrial& c, double ax, double xt0, double xt4, bool any_four,
                                                   bool* need)
{
    double I[5];
    int neg = 0;
    bool bad = false;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const double xt = j == 0 ? xt0 : (j == 4 ? xt4 : ax - G.tn[j]);
        I[j] = lean_node_small<KMAX>(G, L, c, T, xt, any_four, neg, bad);
    }
    const double out = lean_outer(L, I, bad);
    *need = bad | (neg < 0);
    return out;
}

// A trial whose five decision times do not share one regime (common in observed data, where RTs sit a few hundred
// ms after the onset): nodes on both sides of the small-/large-time edge (normalised time ~ 0.15) take the evaluator
// of their own side; a node at or before the onset is exactly 0 (src/pdf.pxi:92-93); a node in (0, xz) is negligible
// and taken as 0 (GridSet::xt_zero, checked against the result at the end as in full_pdf_grid); a node below the
// comfortable range otherwise defers the trial.  Looped over the nodes; the outer level accumulates with the Simpson
// weights.  `mine`: lanes of this class.
template <int KMAX>
__device__ __forceinline__ double lean_trial_mixed(const GridSet<KMAX>& G, const LeanSet& L, const double* __restrict__ T,
                                                   const LeanTrial& c, double ax, bool mine, bool* need)
{
    const int b = c.b;
    int neg = 0;
    bool bad = false, saw_tiny = false;
    double S = 0.0, D = 0.0;
#pragma unroll 1
    for (int j = 0; j < 5; ++j) {
        const double xt = ax - G.tn[j];
        const long long bits = __double_as_longlong(xt);
        const bool pos = mine & (xt > 0.0);
        const bool tiny = pos & (xt < L.xz[b]);
        saw_tiny |= tiny;
        const bool live = pos & !tiny;
        bad |= live & !(bits >= L.x_lo_small[b]);
        const bool sm = live & !bad & (bits < (long long)L.x_edge), lg = live & !bad & !sm;
        const int K = 1 + (int)below(xt, L.x_k[0]) + (int)below(xt, L.x_k[1]) + (int)below(xt, L.x_k[2]);
        const int wK = __reduce_max_sync(0xffffffffu, lg ? K : 1);
        const bool any_four = __any_sync(0xffffffffu, sm & !below(xt, L.x_s4));
        double v = 0.0;
        if (lg) v = lean_node_large<KMAX>(G, L, c, T, xt, exp_lean(G.qk * xt, T), wK, neg, bad);
        if (sm) v = lean_node_small<KMAX>(G, L, c, T, xt, any_four, neg, bad);
        const double ws = (j & 1) ? 4.0 : (j == 2 ? 2.0 : 1.0), wd = (j & 1) ? 4.0 : (j == 2 ? -6.0 : -1.0);
        S = fma(ws, v, S);
        D = fma(wd, v, D);
    }
    bad |= !(fabs(D) <= L.thr_out);
    const double out = fma(D, 1.0 / 15.0, S);
    if (saw_tiny && !(fabs(out) > 1e-150) && !(fabs(out) <= 1e-150 && G.mix_floor > 1e-120)) bad = true;
    *need = bad | (neg < 0);
    return out;
}

// st == 0 (after the snap of src/pdf.pxi:122-123) and sz != 0: full_pdf is the integral over the start point at the
// ONE decision time xt = |rt| - t (src/pdf.pxi:137-141) -- the density tables of gen_rts_from_cdf / gen_cdf_using_pdf
// (src/wfpt.pyx:335,403) are made of such points.  The lean nodes carry k1 = pi / (144 a^2), i.e. the 1/12 of an
// outer level that does not exist here: the value is 12 I''; the stop test is the inner one (thr_in) as it stands.
// All 32 lanes call it (L.ok1 & 1 is set); a lane that needs more than the first level sets *need.
// *n_evals: pdf_sv evaluations of the reference for what was settled here (5; 0 before the onset).
template <int KMAX>
__device__ __forceinline__ double lean_single(const GridSet<KMAX>& G, const LeanSet& L, const double* __restrict__ T,
                                              const LeanTrial& c, double x, bool valid, bool* need, int* n_evals)
{
    const int b = c.b;
    const double ax = fabs(x);
    const double xt = ax - G.t;
    const long long bits = __double_as_longlong(xt);
    const bool before = valid & ((ax - G.t_onset) < 0);      // src/pdf.pxi:111-113: 0, no pdf_sv
    const bool at_onset = valid & !before & (xt <= 0.0);     // five pdf_sv calls that return 0 (src/pdf.pxi:92-93)
    const bool live = valid & (xt > 0.0);
    const bool in_range = live & (bits >= L.x_lo_small[b]) & (xt <= L.xhi);
    const bool sm = in_range & (bits < (long long)L.x_edge) & ((L.ok1 & 2) != 0);
    const bool lg = in_range & !(bits < (long long)L.x_edge);
    bool bad = valid & !(before | at_onset | sm | lg);       // out of range, NaN, a negligible node: complete evaluator
    int neg = 0;
    double v = 0.0;
    if (__any_sync(0xffffffffu, lg)) {
        const int K = 1 + (int)below(xt, L.x_k[0]) + (int)below(xt, L.x_k[1]) + (int)below(xt, L.x_k[2]);
        const int wK = __reduce_max_sync(0xffffffffu, lg ? K : 1);
        bool b1 = false;
        int n1 = 0;
        const double xs = lg ? xt : L.xhi;  // (lanes of other classes: a harmless argument)
        const double vl = lean_node_large<KMAX>(G, L, c, T, xs, exp_lean(G.qk * xs, T), wK, n1, b1);
        if (lg) {
            v = vl;
            bad |= b1;
            neg |= n1;
        }
    }
    if (__any_sync(0xffffffffu, sm)) {
        const bool any_four = __any_sync(0xffffffffu, sm & !below(xt, L.x_s4));
        bool b2 = false;
        int n2 = 0;
        const double xs = sm ? xt : __longlong_as_double(L.x_lo_small[b]);
        const double vs = lean_node_small<KMAX>(G, L, c, T, xs, any_four, n2, b2);
        if (sm) {
            v = vs;
            bad |= b2;
            neg |= n2;
        }
    }
    *need = bad | (neg < 0);
    *n_evals = (valid & !*need & !before) ? 5 : 0;
    return 12.0 * v;
}

// sz == 0 and st == 0 (the simple DDM, with or without sv): full_pdf = pdf_sv(|rt| - t, v, sv, a, z) (src/pdf.pxi:127-128,
// :87-102; sv == 0: pdf, :74-85 -- the same expression with r = 1) -- one series value, one exponential for the drift
// / sv marginal, one more for q (large time) or two for the images (small time).  The node constants are those of the
// lean level (k1 = pi / (144 a^2) carries two Simpson levels that do not exist here: times 144).  All 32 lanes
// call it (L.ok1 & 4 is set); a lane outside the comfortable range or with a negative series value sets *need.
// *n_evals: 1 (0 before the onset).
template <int KMAX>
__device__ __forceinline__ double lean_simple(const GridSet<KMAX>& G, const LeanSet& L, const double* __restrict__ T,
                                              const LeanTrial& c, double x, bool valid, bool* need, int* n_evals)
{
    const int b = c.b;
    const double ax = fabs(x);
    const double xt = ax - G.t;
    const long long bits = __double_as_longlong(xt);
    const bool before = valid & ((ax - G.t_onset) < 0);      // src/pdf.pxi:111-113
    const bool at_onset = valid & !before & (xt <= 0.0);     // pdf_sv(0, ...) = 0 (src/pdf.pxi:92-93)
    const bool live = valid & (xt > 0.0);
    const bool in_range = live & (bits >= L.x_lo_small[b]) & (xt <= L.xhi);
    const bool sm = in_range & (bits < (long long)L.x_edge) & ((L.ok1 & 8) != 0);
    const bool lg = in_range & !(bits < (long long)L.x_edge);
    bool bad = valid & !(before | at_onset | sm | lg);
    int neg = 0;
    double v = 0.0;
    if (__any_sync(0xffffffffu, lg)) {
        const double xs = lg ? xt : L.xhi;
        const int K = 1 + (int)below_pos(xs, L.x_k[0]) + (int)below_pos(xs, L.x_k[1]) + (int)below_pos(xs, L.x_k[2]);
        const int wK = __reduce_max_sync(0xffffffffu, lg ? K : 1);
        const double q = exp_lean(G.qk * xs, T);
        const double q2 = q * q, q3 = q2 * q;
        // series / q at the set's start point: ascending k (src/pdf.pxi:60-62)
        double p = fma(below_pos(xs, L.x_k[0]) ? q3 : 0.0, L.ks[b][1][0], L.ks[b][0][0]);
        if (wK >= 3) {
            const double q4 = q2 * q2, q8 = q4 * q4;
            p = fma(below_pos(xs, L.x_k[1]) ? q8 : 0.0, L.ks[b][2][0], p);
            if (wK >= 4) p = fma(below_pos(xs, L.x_k[2]) ? (q8 * q4) * q3 : 0.0, L.ks[b][3][0], p);
        }
        const double r = rsqrt_fast(fma(G.sv_sq, xs, 1.0));
        const double e = exp_lean(fma(-c.v2h, xs, c.c0h) * (r * r), T);
        if (lg) {
            v = (p * e) * ((r * q) * (144.0 * L.k1));
            neg |= __double2hiint(p);
        }
    }
    if (__any_sync(0xffffffffu, sm)) {
        const double xs = sm ? xt : __longlong_as_double(L.x_lo_small[b]);
        const bool any_four = __any_sync(0xffffffffu, sm & !below(xs, L.x_s4));
        const double rc = rcp_fast(xs);
        const double r = rsqrt_fast(fma(G.sv_sq, xs, 1.0));
        const double u = G.a_sq * rc;
        const double scale = (((u * u) * rsqrt_fast(u)) * r) * (144.0 * L.k1s);
        double C2 = 0.0;
        if (any_four) {  // warp-uniform
            const double C = exp_lean<true>(-L.g4_one * rc, T);
            C2 = below(xs, L.x_s4) ? 0.0 : C * C;
        }
        int n2 = 0;
        const double vs = lean_point_small<KMAX>(G, L, c, T, xs, G.w5[b][0], rc, r * r, scale, C2, n2);
        if (sm) {
            v = vs;
            neg |= n2;
        }
    }
    *need = bad | (neg < 0);
    *n_evals = (valid & !*need & !before) ? 1 : 0;
    return v;
}

// The lean first level for one warp round (all 32 lanes call it; `valid` masks lanes without a trial).  Lean or
// deferred: a lane whose five decision times are all large-time or all small-time and inside the comfortable range
// is evaluated here; a trial before the onset has density exactly 0 and costs the reference no pdf_sv
// (src/pdf.pxi:111-113); any other lane (a node at or before the onset, a trial straddling the small-/large-time
// edge, a failed stop test, ...) sets *need and goes to the deferred-trial pass, whose tree evaluates exactly what
// the reference's recursion reaches -- so the first pass holds no other code for these sets (it was
// instruction-fetch bound on onset-mixed data: profiles/r2_v7_ncu_digest_like_B.txt).  A warp with both classes runs
// both evaluators.  *n_evals = pdf_sv evaluations the reference makes for what was settled here (25 or 0).
template <int KMAX>
__device__ __forceinline__ double lean_first_level(const GridSet<KMAX>& G, const LeanSet& L, const double* __restrict__ T,
                                                   const LeanTrial& c, double x, bool valid, bool* need, int* n_evals)
{
    const int b = c.b;
    const double ax = fabs(x);
    const double xt0 = ax - G.tn[0], xt4 = ax - G.tn[4];
    // x_small: signed compare, a negative or -NaN decision time fails; +NaN fails the second test
    const long long bits4 = __double_as_longlong(xt4);
    const bool large = valid & (bits4 >= L.x_small[b]) & (xt0 <= L.xhi);
    const bool small = valid & !large & (bits4 >= L.x_lo_small[b]) & below(xt0, L.x_edge) & (L.ok_small != 0);
    const bool before = valid & ((ax - G.t_onset) < 0);
    // anything else that starts inside the range: nodes on both sides of the edge, nodes at or before the onset
    const bool mixed = valid & !large & !small & !before & (xt0 <= L.xhi) & (L.ok_small != 0);
    bool nd = valid & !(large | small | mixed | before);
    double d = 0.0;
    if (__any_sync(0xffffffffu, large)) {
        const int K = 1 + (int)below(xt4, L.x_k[0]) + (int)below(xt4, L.x_k[1]) + (int)below(xt4, L.x_k[2]);
        const int wK = __reduce_max_sync(0xffffffffu, large ? K : 1);
        bool n1 = false;
        const double dl = lean_trial_large<KMAX>(G, L, T, c, ax, xt0, xt4, wK, &n1);
        if (large) {
            d = dl;
            nd = n1;
        }
    }
    if (__any_sync(0xffffffffu, small)) {
        const bool any_four = __any_sync(0xffffffffu, small & !below(xt0, L.x_s4));
        bool n2 = false;
        const double ds = lean_trial_small<KMAX>(G, L, T, c, ax, xt0, xt4, any_four, &n2);
        if (small) {
            d = ds;
            nd = n2;
        }
    }
    if (__any_sync(0xffffffffu, mixed)) {
        bool n3 = false;
        const double dm = lean_trial_mixed<KMAX>(G, L, T, c, ax, mixed, &n3);
        if (mixed) {
            d = dm;
            nd = n3;
        }
    }
    *need = nd;
    *n_evals = (valid & !nd & !before) ? 25 : 0;  // 5 decision times x 5 start points, src/integrate.pxi:131-133,187-198
    return d;
}

// Deferred trials, HDDM numerics (n_st == 2): the whole outer recursion of src/integrate.pxi:143-178 for one trial,
// lane-synchronous.  Its decision-time nodes are per-set constants (tn17); a lane evaluates a node only if the
// reference's recursion reaches it (first level m = 0,4,8,12,16; the halves add m = 2,6 / 10,14 when the top stop
// test fails; their quarters add the odd m when a half's test -- same threshold on D'', the tolerance halves with the
// interval, :107-112 -- fails; the quarters return without a test, bottom == 0).  A node is 0 when its decision time
// is not positive (src/pdf.pxi:92-93), else a lean small- or large-time node.  Anything else (a node below the
// comfortable range, an inner stop test that fails, a negative series value, NaN) sets *fail: the trial goes on to
// the complete evaluator.  *n_evals = pdf_sv evaluations made (5 per node + those of inner refinements).
template <int KMAX>
__device__ __noinline__ double lean_tree(const GridSet<KMAX>& G, const LeanSet& L, const double* __restrict__ T,
                                         double x, const LeanTrial& c, bool active, bool* fail, int* n_evals)
{
    const int b = c.b;
    const double ax = fabs(x);
    bool bad = !(ax == ax) || !(ax - G.t_onset >= 0);  // NaN; before the onset (full_pdf returns 0 without a node)
    DBG(0, active); DBG(2, active & bad);
    int neg = 0, count = 0, extra = 0;
    bool saw_tiny = false;
    // value of node m for the lanes in `want` (others keep 0); warp-convergent
    auto node = [&](int m, bool want) {
        const double xt = ax - L.tn17[m];
        const bool pos = xt > 0.0;
        const long long bits = __double_as_longlong(xt);
        // a decision time in (0, xz): the node is negligible (GridSet::xt_zero) and taken as 0; checked at the end
        const bool tiny = want & pos & (xt < L.xz[b]);
        saw_tiny |= tiny;
        const bool small = want & pos & !tiny & (bits < (long long)L.x_edge);
        const bool large = want & pos & !tiny & !small;
        DBG(3, small & !bad & !((bits >= L.x_lo_small[b]) & (L.ok_small != 0)));
        DBG(4, large & !bad & !(xt <= L.xhi));
        bad |= small & !((bits >= L.x_lo_small[b]) & (L.ok_small != 0));
        bad |= large & !(xt <= L.xhi);
        const bool do_small = small & !bad, do_large = large & !bad;
        const bool bad_before = bad;
        const int K = 1 + (int)below(xt, L.x_k[0]) + (int)below(xt, L.x_k[1]) + (int)below(xt, L.x_k[2]);
        const int wK = __reduce_max_sync(0xffffffffu, do_large ? K : 1);
        const bool any_four = __any_sync(0xffffffffu, do_small & !below(xt, L.x_s4));
        double val = 0.0;
        if (do_large) val = lean_node_large<KMAX>(G, L, c, T, xt, exp_lean(G.qk * xt, T), wK, neg, bad);
        if (do_small) {
            bool pass = true;
            double f5[5];
            val = lean_node_small<KMAX>(G, L, c, T, xt, any_four, neg, bad, &pass, f5);
            if (!pass) val = lean_refine_small<KMAX>(G, L, c, T, xt, f5, neg, &extra);  // (NaN values fail every test)
        }
        DBG(5, do_large & bad & !bad_before); DBG(6, do_small & bad & !bad_before); DBG(7, want);
        count += want ? 1 : 0;
        return val;
    };
    // Simpson on five equally spaced values: S'' + D'' / 15 and whether the stop test passes
    auto simpson = [&](double y0, double y1, double y2, double y3, double y4, bool* pass) {
        const double s04 = y0 + y4, s13 = y1 + y3;
        const double S = fma(2.0, y2, fma(4.0, s13, s04));
        const double D = fma(-6.0, y2, fma(4.0, s13, -s04));
        *pass = fabs(D) <= L.thr_out;
        return fma(D, 1.0 / 15.0, S);
    };
    double I[17];  // (indexed in a loop: lives in local memory -- 17 words against ~100 instructions per node)
#pragma unroll
    for (int m = 0; m < 17; ++m) I[m] = 0.0;
    active &= !bad;
    // ONE loop over the 17 nodes in level order, so that the node evaluator is inlined once (the deferred-trial
    // pass is instruction-fetch bound: profiles/r2_v7_ncu_digest_like_B_second.txt)
    bool refine = false, ql = false, qr = false;
    double out = 0.0, left = 0.0, right = 0.0;
#pragma unroll 1
    for (int s = 0; s < 17; ++s) {
        if (s == 5) {  // first level done: the top stop test
            bool top_ok;
            out = simpson(I[0], I[4], I[8], I[12], I[16], &top_ok);
            refine = active & !top_ok & !bad;  // (a NaN D fails the test, as in the reference)
            if (!__any_sync(0xffffffffu, refine)) break;
        }
        if (s == 9) {  // halves done: their stop tests (same threshold on D'': the tolerance halves with the interval)
            bool left_ok, right_ok;
            left = 0.5 * simpson(I[0], I[2], I[4], I[6], I[8], &left_ok);
            right = 0.5 * simpson(I[8], I[10], I[12], I[14], I[16], &right_ok);
            ql = refine & !left_ok & !bad;
            qr = refine & !right_ok & !bad;
            if (!__any_sync(0xffffffffu, ql | qr)) break;
        }
        // level order: 0 4 8 12 16 | 2 6 10 14 | 1 3 5 7 9 11 13 15
        const int m = s < 5 ? 4 * s : (s < 9 ? 4 * (s - 5) + 2 : 2 * (s - 9) + 1);
        const bool want = s < 5 ? active : (s < 9 ? refine : (m < 8 ? ql : qr));
        if (!__any_sync(0xffffffffu, want)) continue;
        I[m] = node(m, want);
    }
    {
        bool dummy;
        if (ql) left = 0.25 * simpson(I[0], I[1], I[2], I[3], I[4], &dummy) +
                       0.25 * simpson(I[4], I[5], I[6], I[7], I[8], &dummy);
        if (qr) right = 0.25 * simpson(I[8], I[9], I[10], I[11], I[12], &dummy) +
                        0.25 * simpson(I[12], I[13], I[14], I[15], I[16], &dummy);
        if (refine) out = left + right;
    }
    // negligible nodes were taken as 0: fine unless the whole result is that small too and no outlier term swamps it
    if (saw_tiny && !(fabs(out) > 1e-150) && !(fabs(out) <= 1e-150 && G.mix_floor > 1e-120)) bad = true;
    if (!(out == out)) bad = true;  // NaN: let the complete evaluator reproduce the reference's NaN semantics
    DBG(1, active & (bad | (neg < 0))); DBG(8, active & (neg < 0)); DBG(9, refine);
    *fail = bad | (neg < 0);
    *n_evals = 5 * count + extra;
    return out;
}

}  // namespace hddm
