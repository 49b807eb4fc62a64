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

constexpr int kExpBigBits = 11;
constexpr int kExpBig = 1 << kExpBigBits;
// 2^(i/2048), filled once per process by the host (launch_like -> ensure_exp_table) with correctly rounded values.
__device__ double g_exp2_big[kExpBig];

// exp(x), |x| <= 128 (callers guarantee it): x = n ln2/2048 + r, |r| <= ln2/4096 = 1.7e-4;
// exp(x) = 2^(n >> 11) T[n & 2047] (1 + r + r^2/2 + r^3/6), truncation r^4/24 < 4e-17, argument reduction
// with the single constant ln2/2048 (its rounding error times |n| <= 3.8e5 stays below 1.5e-14 |x| / 128).
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
    return __hiloint2double(__double2hiint(m) + ((n >> kExpBigBits) << 20), __double2loint(m));
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
    if (v != 0.0 && room / (v * v) < xhi) xhi = room / (v * v);
    if (G.xt_hi < xhi) xhi = G.xt_hi;
    L.xhi = xhi;
    L.ok = G.ok && G.support && drift_uniform && mono && G.sz != 0 && G.st != 0 && nm.n_sz > 0 && nm.n_st > 0 &&
           nm.simps_err >= 0.0 && asw * asw <= 50.0 && xhi > 0.0 && isfinite(v) && isfinite(L.k1) && L.k1 > 0.0 ? 1 : 0;
    L.pad = 0;
}

__device__ __forceinline__ bool below(double xt, unsigned long long edge)
{
    return (unsigned long long)__double_as_longlong(xt) < edge;
}

// One trial, all five decision times, at most WK large-time terms per node (WK is the warp's maximum).
// b: boundary (0 lower / 1 upper); ax = |rt|; xt0 = ax - tn[0] and xt4 = ax - tn[4] have passed the range checks.
// Returns full_pdf; *need is set when a stop test fails or a series value is negative (the trial is then
// re-evaluated by the complete evaluator, and the return value is ignored).
template <int KMAX, int WK>
__device__ __forceinline__ double lean_trial(const GridSet<KMAX>& G, const LeanSet& L, const double* __restrict__ T,
                                             int b, double ax, double xt0, double xt4, bool* need)
{
    const double q0 = exp_lean(G.qk * xt0, T);
    const double c0h = L.c0h[b], dc0h = L.dc0h[b], d2ch = L.d2ch[b];
    double I[5];
    int neg = 0;
    bool bad = false;
#pragma unroll
    for (int j = 0; j < 5; ++j) {
        const double xt = j == 0 ? xt0 : (j == 4 ? xt4 : ax - G.tn[j]);
        const double q = j == 0 ? q0 : q0 * G.rho[j];
        // series / q: sin(pi w_i) + sum_{k>=2} q^(k^2-1) k sin(k pi w_i), ascending k (src/pdf.pxi:60-62)
        double p[5], s[5];
        load5(L.ks[b][0], p);
        if (WK >= 2) {
            const double q2 = q * q, q3 = q2 * q;
            const double e2 = below(xt, L.x_k[0]) ? q3 : 0.0;
            load5(L.ks[b][1], s);
#pragma unroll
            for (int i = 0; i < 5; ++i) p[i] = fma(e2, s[i], p[i]);
            if (WK >= 3) {
                const double q4 = q2 * q2, q8 = q4 * q4;
                const double e3 = below(xt, L.x_k[1]) ? q8 : 0.0;
                load5(L.ks[b][2], s);
#pragma unroll
                for (int i = 0; i < 5; ++i) p[i] = fma(e3, s[i], p[i]);
                if (WK >= 4) {
                    const double q15 = (q8 * q4) * q3;
                    const double e4 = below(xt, L.x_k[2]) ? q15 : 0.0;
                    load5(L.ks[b][3], s);
#pragma unroll
                    for (int i = 0; i < 5; ++i) p[i] = fma(e4, s[i], p[i]);
                }
            }
        }
        neg |= (__double2hiint(p[0]) | __double2hiint(p[1])) | (__double2hiint(p[2]) | __double2hiint(p[3])) |
               __double2hiint(p[4]);
        // sv marginal (src/pdf.pxi:102): exp(c(w_i) d - v^2 xt d) r / a^2, d = 1 / (2 (sv^2 xt + 1)) = r^2 / 2; the
        // numerator is quadratic in i, so the five exponentials are a second-order geometric chain from three
        const double r = rsqrt_fast(fma(G.sv_sq, xt, 1.0));
        const double dd = r * r;
        const double E0 = exp_lean(fma(-L.v2h, xt, c0h) * dd, T) * ((r * q) * L.k1);
        const double g0 = exp_lean(dc0h * dd, T);
        const double r2 = exp_lean(d2ch * dd, T);
        const double E1 = E0 * g0, g1 = g0 * r2;
        const double E2 = E1 * g1, g2 = g1 * r2;
        const double E3 = E2 * g2, g3 = g2 * r2;
        const double E4 = E3 * g3;
        // inner Simpson level over the five start points (src/integrate.pxi:114-141, :99-106)
        const double s04 = fma(p[4], E4, p[0] * E0);
        const double s13 = fma(p[3], E3, p[1] * E1);
        const double f2 = p[2] * E2;
        const double S = fma(2.0, f2, fma(4.0, s13, s04));
        const double D = fma(-6.0, f2, fma(4.0, s13, -s04));
        bad |= !(fabs(D) <= L.thr_in);
        I[j] = fma(D, 1.0 / 15.0, S);
    }
    // outer Simpson level over the five decision times (src/integrate.pxi:143-178)
    const double s04 = I[0] + I[4], s13 = I[1] + I[3];
    const double S = fma(2.0, I[2], fma(4.0, s13, s04));
    const double D = fma(-6.0, I[2], fma(4.0, s13, -s04));
    bad |= !(fabs(D) <= L.thr_out);
    *need = bad | (neg < 0);
    return fma(D, 1.0 / 15.0, S);
}

}  // namespace hddm
