// wfpt_grid.cuh -- v3 evaluator of full_pdf for the headline case: every parameter except the drift
// is uniform over the CTA's parameter set and integration is adaptive.
//
// The reference's recursion (src/integrate.pxi:114-206) always evaluates its first level: 5 decision
// times t' (lb, ub, c, d, e of the outer Simpson rule) x 5 start points w, and its stop tests almost
// always accept that level (mean 24.9 pdf_sv per eval).  This file makes that 5 x 5
// grid cheap; everything deeper, and every trial outside the "comfortable" range defined below, goes
// through the v2 walk of wfpt_fast.cuh unchanged.
//
// Per parameter set (GridSet, shared memory, built once per CTA):
//   * start-point tables of wfpt_fast.cuh (abscissae by the reference's midpoint recursion, sin(k pi w));
//   * the five first-level decision-time offsets tn[j] (same midpoint recursion, src/integrate.pxi:153-155,
//     :187-198) and rho[j] = exp(pi^2 (tn[j]-tn[0]) / (2 a^2)): the large-time q = exp(-pi^2 tt/2) of
//     node j is q_0 * rho[j], one exp per trial instead of five;
//   * the break points of the (branch, K) decision mapped from normalised time to decision time:
//     xthr[i] = the smallest double X with fl(X / a^2) >= thr[i], so `x_t >= xthr[i]` (an integer
//     compare of bit patterns, off the FP64 pipe) decides exactly as the reference's tt = x/a^2 does;
//   * S w_i^2 and 2 a w_i of the sv-marginal exponent at the five start points.
// Per (time node, start point): K FMAs (large-time) or K table-driven exps (small-time), one
// table-driven exp for the sv marginal.  exp is a branch-free 64-entry-table / degree-5 evaluation
// (10 FP64 instructions, < 1 ulp + 2e-17) -- the libdevice exp was the dominant cost of v2.
// DECISIONS (support test, branch, K, every Simpson stop test) remain the reference's.
#pragma once
#include "wfpt_fast.cuh"

namespace hddm {

constexpr int kExpTab = 64;
constexpr int kThrWin = 8;  // thresholds of the series decision that can fall inside the comfortable range

// 2^(i/64), correctly rounded.
__device__ const double g_exp2_tab[kExpTab] = {
    0x1.0000000000000p+0, 0x1.02c9a3e778061p+0, 0x1.059b0d3158574p+0, 0x1.0874518759bc8p+0,
    0x1.0b5586cf9890fp+0, 0x1.0e3ec32d3d1a2p+0, 0x1.11301d0125b51p+0, 0x1.1429aaea92de0p+0,
    0x1.172b83c7d517bp+0, 0x1.1a35beb6fcb75p+0, 0x1.1d4873168b9aap+0, 0x1.2063b88628cd6p+0,
    0x1.2387a6e756238p+0, 0x1.26b4565e27cddp+0, 0x1.29e9df51fdee1p+0, 0x1.2d285a6e4030bp+0,
    0x1.306fe0a31b715p+0, 0x1.33c08b26416ffp+0, 0x1.371a7373aa9cbp+0, 0x1.3a7db34e59ff7p+0,
    0x1.3dea64c123422p+0, 0x1.4160a21f72e2ap+0, 0x1.44e086061892dp+0, 0x1.486a2b5c13cd0p+0,
    0x1.4bfdad5362a27p+0, 0x1.4f9b2769d2ca7p+0, 0x1.5342b569d4f82p+0, 0x1.56f4736b527dap+0,
    0x1.5ab07dd485429p+0, 0x1.5e76f15ad2148p+0, 0x1.6247eb03a5585p+0, 0x1.6623882552225p+0,
    0x1.6a09e667f3bcdp+0, 0x1.6dfb23c651a2fp+0, 0x1.71f75e8ec5f74p+0, 0x1.75feb564267c9p+0,
    0x1.7a11473eb0187p+0, 0x1.7e2f336cf4e62p+0, 0x1.82589994cce13p+0, 0x1.868d99b4492edp+0,
    0x1.8ace5422aa0dbp+0, 0x1.8f1ae99157736p+0, 0x1.93737b0cdc5e5p+0, 0x1.97d829fde4e50p+0,
    0x1.9c49182a3f090p+0, 0x1.a0c667b5de565p+0, 0x1.a5503b23e255dp+0, 0x1.a9e6b5579fdbfp+0,
    0x1.ae89f995ad3adp+0, 0x1.b33a2b84f15fbp+0, 0x1.b7f76f2fb5e47p+0, 0x1.bcc1e904bc1d2p+0,
    0x1.c199bdd85529cp+0, 0x1.c67f12e57d14bp+0, 0x1.cb720dcef9069p+0, 0x1.d072d4a07897cp+0,
    0x1.d5818dcfba487p+0, 0x1.da9e603db3285p+0, 0x1.dfc97337b9b5fp+0, 0x1.e502ee78b3ff6p+0,
    0x1.ea4afa2a490dap+0, 0x1.efa1bee615a27p+0, 0x1.f50765b6e4540p+0, 0x1.fa7c1819e90d8p+0};

__constant__ double g_exp_c[6] = {0x1.71547652b82fep+6, -0x1.62e42fefa39efp-7, -0x1.abc9e3b39803fp-62,
                                  1.0 / 120.0, 1.0 / 24.0, 1.0 / 6.0};

// exp(x) for -700 <= x <= 700 (CLAMP: -2e7 <= x <= 700; results below 2^-1021 come out as exactly 0,
// used only where such a term is added to one that is larger by tens of orders of magnitude).
// x = (64 n' + i) ln2/64 + r, |r| <= ln2/128: exp(x) = 2^n' * T[i] * (1 + r + r^2/2 + ... + r^5/120).
template <bool CLAMP>
__device__ __forceinline__ double exp_tab(double x, const double* __restrict__ T)
{
    const double magic = 6755399441055744.0;  // 1.5 * 2^52: adding it leaves rint(.) in the low word
    const double t = fma(x, g_exp_c[0], magic);
    const int n = __double2loint(t);
    const double nd = t - magic;
    double r = fma(nd, g_exp_c[1], x);
    r = fma(nd, g_exp_c[2], r);
    const double r2 = r * r;
    double c = fma(r, g_exp_c[3], g_exp_c[4]);
    c = fma(c, r, g_exp_c[5]);
    c = fma(c, r, 0.5);
    const double p = fma(r2, c, r);
    const double tv = T[n & (kExpTab - 1)];
    const double m = fma(tv, p, tv);
    const int e = n >> 6;
    int hi = __double2hiint(m) + (e << 20), lo = __double2loint(m);
    if (CLAMP && e < -1021) hi = lo = 0;  // below the normal range: exactly 0
    return __hiloint2double(hi, lo);
}

// 1/sqrt(y), y normal and positive: MUFU seed (~2^-20) + one cubic step (error ~ 5/16 e^3 < 1e-17).
__device__ __forceinline__ double rsqrt_fast(double y)
{
    double s;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(s) : "d"(y));
    const double e = fma(-(y * s), s, 1.0);
    const double c = fma(0.375, e, 0.5);
    return fma(s * e, c, s);
}

// 1/y, y normal and positive: MUFU seed + one cubic step.
__device__ __forceinline__ double rcp_fast(double y)
{
    double s;
    asm("rcp.approx.ftz.f64 %0, %1;" : "=d"(s) : "d"(y));
    const double e = fma(-y, s, 1.0);
    return fma(fma(e, e, e), s, s);
}

// Per-set constants of the lean first level (wfpt_lean.cuh); part of GridSet so that it is built once per set
// and copied with it.
struct alignas(16) LeanSet {
    double ks[2][4][6];   // k sin(k pi w_i), k = 1..4, at the five first-level start points (padded)
    double c0h[2], dc0h[2], d2ch[2];  // halves of c(w_0), c(w_1) - c(w_0) and the second difference (GridSet::sw0..d2c)
    double xhi;           // largest decision time the lean level takes: exponent arguments stay inside [-128, 128]
    double v2h;           // v^2 / 2
    double k1;            // pi / (144 a^2)
    double thr_in, thr_out;  // 15 simps_err / 12 and 15 simps_err
    // bit patterns of decision-time break points: a node is large-time and inside the comfortable range of
    // its boundary iff xt >= x_small[b] (compared as SIGNED integers: a negative xt fails); it has >= k terms iff
    // xt < x_k[k - 2] (k = 2, 3, 4)
    long long x_small[2];
    unsigned long long x_k[3];
    int ok, ok_small;
    // Small-time nodes (src/pdf.pxi:51-57) with 3 or 4 terms: exp(-(w_i + 2k)^2 / (2 tt)) = A_i R_{k,i} (see grid_node);
    // every exponent is a per-set constant times 1 / xt.  g_* = a^2 / 2 times w_0^2, w_1^2 - w_0^2, the second
    // difference of w_i^2; g4_* = 2 a^2 times h_w, 1, 1 - w_4, w_0 + 1.
    double g_w0sq[2], g_dw1[2], g_d2w[2], g4_hw, g4_one, g4_omw4[2], g4_w0p1[2];
    double k1s;            // 1 / (144 a^2 sqrt(2 pi))
    // a node is small-time iff xt < x_edge (the bit pattern behind x_small); it has 4 terms (k = -1..2) iff
    // xt >= x_s4, else 3 (k = -1..1); the lean small-time node takes xt >= x_lo_small[b] (signed compare)
    unsigned long long x_edge, x_s4;
    long long x_lo_small[2];
    // decision-time nodes of the outer recursion down to its third level (n_st = 2: lb + m st / 16, m = 0..16, by the
    // reference's midpoint rule, src/integrate.pxi:153-155): the deferred-trial tree of wfpt_lean.cuh
    double tn17[17];
    double v;              // the set's drift (boundary flip applied per trial)
    int uniform;           // the drift-dependent constants above belong to the set (else: per trial, lean_begin_v)
    int ok1;               // st == 0 (one decision time, src/pdf.pxi:137-141): bit 0 lean_single applies, bit 1 to small-time nodes too
    double xz[2];          // decision times in (0, xz[b]) are negligible nodes (GridSet::xt_zero; 0: none are)
    double pad17;
};

template <int KMAX>
struct alignas(16) GridSet {
    double s5[2][KMAX][6];  // sin(k pi w_i) at the five first-level start points (table nodes 4 i), padded
    double w5[2][6];        // those start points
    WTable<KMAX> wt[2];
    double etab[kExpTab];
    double tn[5];          // decision-time nodes of the first outer level, ascending: lb, d, c, e, ub
    double rho[5];         // exp(pi^2 (tn[j] - tn[0]) / (2 a^2))
    // sv-marginal exponent numerator c(w) = (a sv w)^2 - 2 a v w at the start points, as a function of
    // the drift: c_0 = sw0 - aw0 v, c_1 - c_0 = dsw - daw v, second difference d2c (free of v)
    double sw0[2], aw0[2], dsw[2], daw[2], d2c[2];
    // small-time chain (see grid_node): w_0^2, w_1^2 - w_0^2, second difference of w_i^2, w_0 + 1,
    // 1 - w_4 per boundary; hw = w_1 - w_0; chain_ok: start points well inside (0, 1)
    double w0sq[2], dw1[2], d2w[2], w0p1[2], omw4[2], hw;
    int chain_ok[2];
    unsigned long long xwin[kThrWin];  // bit patterns of xthr inside the window, padded with +inf
    int code[kMaxRegions];
    int region0;           // thresholds below the window (always passed)
    int ok;                // 0: the grid evaluator cannot be used for this set
    int support;           // the set-level part of the support test of src/pdf.pxi:111-113 passed
    double t_onset;        // t - st/2 with the raw st (the per-trial part of that test)
    double xt_lo[2], xt_hi;  // comfortable decision-time range per boundary, see build_gridset
    double xt_zero[2];       // below this decision time every start point's density is < 1e-230, see build_gridset
    double mix_floor;        // w_outlier * p_outlier the caller adds to every density (0: none), see full_pdf_grid
    double a, sv, t, st, sz;  // st, sz after the 1e-3 snap of src/pdf.pxi:122-125
    double a_sq, inv_a_sq, sv_sq, qk;  // qk = -pi^2 / (2 a^2)
    double inv_zt, inv_st; // 1/(ub_z - lb_z) (1 when sz == 0), 1/(ub_t - lb_t)
    double inv_a_sq_zt;    // inv_a_sq * inv_zt
    double zw, zw6, zw12;  // ub_z - lb_z, and that / 6, / 12 (src/integrate.pxi:99-100,137)
    double tw, tw6, tw12;  // ub_t - lb_t, / 6, / 12
    double raw[7];         // the set's v, sv, a, z, sz, t, st as given (for the out-of-range fallback)
    LeanSet lean;          // lean.ok == 0 unless the builder was asked for it (first pass of long segments)
};

// Smallest double X with fl(X / a_sq) >= thr (thr and thr * a_sq normal and positive).
__device__ __forceinline__ double xt_threshold(double thr, double a_sq)
{
    double X = thr * a_sq;
    for (int it = 0; it < 8 && X / a_sq >= thr; ++it) X = __longlong_as_double(__double_as_longlong(X) - 1);
    for (int it = 0; it < 16 && X / a_sq < thr; ++it) X = __longlong_as_double(__double_as_longlong(X) + 1);
    return X;
}

// Cooperative build (a CTA of >= 256 threads, or one warp: every phase runs over 256 "virtual thread ids", one
// per thread of a CTA, eight per lane of a warp -- the same arithmetic either way, so the tables are bit-identical).
// par = v, sv, a, z, sz, t, st of the set (v unused).
template <int KMAX, class Group = BlockGroup>
__device__ __forceinline__ void build_gridset(GridSet<KMAX>& G, const double par[7], const SeriesTable& tab,
                                              double mix_floor = 0.0, Group grp = Group())
{
    build_wtables<KMAX>(G.wt, par[3], par[4], grp);  // ends with a barrier
  for (int vj = 0; vj < virtual_rounds(grp); ++vj) {
    const int tid = virtual_tid(grp, grp.tid(), vj);
    const double a = par[2], sv = par[1];
    const double a_sq = a * a;
    const double st = par[6] < 1e-3 ? 0.0 : par[6];
    const double sz = par[4] < 1e-3 ? 0.0 : par[4];
    const double tt_lo = 1.0 / 1048576.0, tt_hi = 128.0;
    if (tid < kExpTab) G.etab[tid] = g_exp2_tab[tid];
    if (tid >= 72 && tid < 79) G.raw[tid - 72] = par[tid - 72];
    if (tid == 64) {
        const double lo = par[5] - st / 2., hi = par[5] + st / 2.;
        const double c = (lo + hi) / 2.;
        G.tn[0] = lo; G.tn[4] = hi; G.tn[2] = c;
        G.mix_floor = mix_floor;
        G.tn[1] = (lo + c) / 2.; G.tn[3] = (c + hi) / 2.;
        G.a = a; G.sv = sv; G.t = par[5]; G.st = st; G.sz = sz;
        G.a_sq = a_sq; G.inv_a_sq = 1.0 / a_sq; G.sv_sq = sv * sv; G.qk = -kPiSq / (2.0 * a_sq);
        G.zw = G.wt[0].w[16] - G.wt[0].w[0];
        G.zw6 = G.zw / 6.0; G.zw12 = G.zw / 12.0;
        G.tw = hi - lo; G.tw6 = G.tw / 6.0; G.tw12 = G.tw / 12.0;
        G.inv_zt = sz == 0 ? 1.0 : 1.0 / G.zw;
        G.inv_a_sq_zt = (1.0 / a_sq) * G.inv_zt;
        G.inv_st = st == 0 ? 1.0 : 1.0 / (hi - lo);
        // Comfortable range of normalised time: [max(2^-20, w_0^2 / 1300), 128].  Lower end: the largest
        // small-time term exp(-w_0^2 / (2 tt)) stays above e^-650 (every term that underflows is then < 1e-25 of
        // it) and exp arguments stay above -2e7; upper end: q = exp(-pi^2 tt / 2) stays normal.
        for (int b = 0; b < 2; ++b) {
            const double w0 = G.wt[b].w[0];
            const double lo_b = (w0 * w0) / 1300.0;
            G.xt_lo[b] = (lo_b > tt_lo ? lo_b : tt_lo) * a_sq;
        }
        G.xt_hi = tt_hi * a_sq;
        // Below tt = w_0^2 / 1300 every start point has exp(-w^2 / (2 tt)) <= e^-650; with the sv-marginal
        // exponent bounded by 100 + 100 (checked per set here and per trial in full_pdf_grid) the node's
        // density is < 1e-230: when the trial's result is > 1e-150 such a node is exactly as good as 0.
        {
            const double wm = G.wt[0].w[16] > G.wt[1].w[16] ? G.wt[0].w[16] : G.wt[1].w[16];
            const double asw_ = (a * sv) * wm;
            for (int b = 0; b < 2; ++b) {
                const double w0 = G.wt[b].w[0];
                G.xt_zero[b] = (asw_ * asw_ <= 200.0) ? ((w0 * w0) / 1300.0) * a_sq : 0.0;
            }
        }
        // window of thresholds inside (tt_lo, tt_hi]
        int r0 = 0, nwin = 0;
        for (int i = 0; i < tab.n_thr; ++i) {
            if (tab.thr[i] <= tt_lo) r0++;
            else if (tab.thr[i] <= tt_hi) nwin++;
        }
        G.region0 = r0;
        {
            const double z = par[3], szr = par[4], t = par[5], str = par[6];
            G.support = !((z < 0) || (z > 1) || (a < 0) || (t < 0) || (str < 0) || (sv < 0) || (szr < 0) ||
                          (szr > 1) || (z + szr / 2. > 1) || (z - szr / 2. < 0) || (t - str / 2. < 0));
            G.t_onset = t - str / 2.;
        }
        // a in a sane range keeps every product below normal and finite
        // and the positive part of the sv-marginal exponent, (a sv w)^2 / (2 sv^2 x + 2) <= (a sv w_4)^2 / 2, stays small
        const double wmax = G.wt[0].w[16] > G.wt[1].w[16] ? G.wt[0].w[16] : G.wt[1].w[16];
        const double asw = (a * sv) * wmax;
        G.ok = (nwin <= kThrWin - 1) && (a_sq >= 1e-100) && (a_sq <= 1e100) && (tab.n_thr >= 0) &&
               (asw * asw <= 600.0) ? 1 : 0;
    }
    if (tid >= 96 && tid < 96 + kMaxRegions) G.code[tid - 96] = tid - 96 <= tab.n_thr ? tab.code[tid - 96] : 0;
    if (tid >= 80 && tid < 82) {
        const int b = tid - 80;
        double sw[3], aw[3];
        for (int i = 0; i < 3; ++i) {
            const double w = G.wt[b].w[4 * i];
            const double asw = (a * sv) * w;
            sw[i] = asw * asw;
            aw[i] = (2.0 * a) * w;
        }
        G.sw0[b] = sw[0]; G.aw0[b] = aw[0];
        G.dsw[b] = sw[1] - sw[0]; G.daw[b] = aw[1] - aw[0];
        G.d2c[b] = (sw[2] - sw[1]) - (sw[1] - sw[0]);
        for (int i = 0; i < 6; ++i) G.w5[b][i] = i < 5 ? G.wt[b].w[4 * i] : 0.0;
        {
            const double w0 = G.wt[b].w[0], w1 = G.wt[b].w[4], w2 = G.wt[b].w[8], w4 = G.wt[b].w[16];
            G.w0sq[b] = w0 * w0;
            G.dw1[b] = (w1 - w0) * (w1 + w0);
            G.d2w[b] = (w2 - w1) * (w2 + w1) - (w1 - w0) * (w1 + w0);
            G.w0p1[b] = w0 + 1.0;
            G.omw4[b] = 1.0 - w4;
            if (b == 0) G.hw = w1 - w0;
            G.chain_ok[b] = (w0 >= 0.01) && (w4 <= 0.99) ? 1 : 0;
        }
    }
    if (tid >= 160 && tid < 160 + 2 * KMAX * 6) {
        const int r = tid - 160, b = r / (KMAX * 6), k = (r / 6) % KMAX, i = r % 6;
        G.s5[b][k][i] = i < 5 ? G.wt[b].s[k][4 * i] : 0.0;
    }
  }
    grp.sync();
  for (int vj = 0; vj < virtual_rounds(grp); ++vj) {
    const int tid = virtual_tid(grp, grp.tid(), vj);
    if (tid < 5) G.rho[tid] = exp(-G.qk * (G.tn[tid] - G.tn[0]));
    if (tid >= 32 && tid < 32 + kThrWin) {
        const int k = tid - 32, i = G.region0 + k;
        double X = __longlong_as_double(0x7ff0000000000000LL);
        if (G.ok && i < tab.n_thr && tab.thr[i] <= 128.0) X = xt_threshold(tab.thr[i], par[2] * par[2]);
        G.xwin[k] = (unsigned long long)__double_as_longlong(X);
    }
  }
    grp.sync();
}

// (branch, K) code for a positive decision time inside the comfortable range: three-step binary
// search for #{k : xt >= xwin[k]} over the 7 usable window entries (xwin[7] is always +inf).
template <int KMAX>
__device__ __forceinline__ int grid_code(const GridSet<KMAX>& G, double xt)
{
    const unsigned long long bits = (unsigned long long)__double_as_longlong(xt);
    int r = bits >= G.xwin[3] ? 4 : 0;
    r += bits >= G.xwin[r + 1] ? 2 : 0;
    r += bits >= G.xwin[r] ? 1 : 0;
    return G.code[G.region0 + r];
}

// Per-trial constants of the grid evaluator (after the boundary flip).
struct GridTrial {
    double v, v_sq;
    double c0, dc0;  // c(w_0) and c(w_1) - c(w_0) for this trial's drift
    int b;
};

__device__ __forceinline__ void load5(const double* __restrict__ p, double v[5])
{
    const double2 a = *reinterpret_cast<const double2*>(p);
    const double2 b = *reinterpret_cast<const double2*>(p + 2);
    v[0] = a.x; v[1] = a.y; v[2] = b.x; v[3] = b.y; v[4] = p[4];
}

// pdf_sv(xt, v, sv, a, w_i) / ZT at the NW first-level start points (NW = 5: table nodes 0,4,..,16;
// NW = 1: the set's own start point) of one decision time xt inside the comfortable range.
// `code` is the series decision at xt, `q` the large-time exp(-pi^2 tt / 2) (unused for small-time).
template <int KMAX, int NW>
__device__ __forceinline__ void grid_node(const GridSet<KMAX>& G, const GridTrial& c, double xt, int code,
                                          double q, double f[NW])
{
    double p[NW];
    double post;
    auto sines = [&](int k, double sv[NW]) {  // sin(k pi w_i), k = 1..KMAX
        if (NW == 5) load5(G.s5[c.b][k - 1], sv);
        else sv[0] = G.s5[c.b][k - 1][2];
    };
    if (code > 0) {  // small-time, src/pdf.pxi:51-57
        double w[NW];
        if (NW == 5) load5(G.w5[c.b], w);
        else w[0] = G.w5[c.b][2];
#pragma unroll
        for (int i = 0; i < NW; ++i) p[i] = 0.0;
        const double inv_tt = G.a_sq * rcp_fast(xt);
        const double arg = 0.5 * inv_tt;
        const int K = code;
        if (NW == 5 && K <= 5 && G.chain_ok[c.b]) {
            // exp(-g (w_i + 2k)^2) = A_i R_{k,i} with A_i = exp(-g w_i^2) and R_{k,i} = exp(-4 g k (w_i + k)) <= 1.
            // The w_i are equally spaced, so A_i is a quadratic chain (3 exps), R_{1,i} and R_{-1,i} are geometric
            // in i with ratio eta = exp(-4 g h) (3 exps), R_{+-2,i} = R_{+-1,i}^2 exp(-8 g) (1 exp): 7 exps instead
            // of 5 K.  A side whose largest ratio is below e^-60 cannot change the sum (it is < 1e-23 of the
            // k = 0 term) and is skipped.
            const int b = c.b;
            double A = exp_tab<false>(-arg * G.w0sq[b], G.etab);
            double ga = exp_tab<true>(-arg * G.dw1[b], G.etab);
            const double de = exp_tab<true>(-arg * G.d2w[b], G.etab);
            const double g4 = 4.0 * arg;
            const bool pos = g4 * G.w0p1[b] <= 60.0;            // k = 1 (and 2)
            const bool neg = K >= 3 && g4 * G.omw4[b] <= 60.0;  // k = -1 (and -2)
            double t[5];
#pragma unroll
            for (int i = 0; i < 5; ++i) t[i] = w[i];
            if (pos || neg) {
                const double eta = exp_tab<true>(-g4 * G.hw, G.etab);
                double C2 = 0.0;
                if (K >= 4) {
                    const double C = exp_tab<true>(-g4, G.etab);
                    C2 = C * C;
                }
                if (neg) {  // the reference adds k ascending; negative k first
                    double R = exp_tab<true>(-g4 * G.omw4[b], G.etab);
#pragma unroll
                    for (int i = 4; i >= 0; --i) {
                        double u = (w[i] - 2.0) * R;
                        if (K >= 5) u = fma(w[i] - 4.0, (R * R) * C2, u);
                        t[i] = u + t[i];
                        if (i > 0) R *= eta;
                    }
                }
                if (pos) {
                    double R = exp_tab<true>(-g4 * G.w0p1[b], G.etab);
#pragma unroll
                    for (int i = 0; i < 5; ++i) {
                        t[i] = fma(w[i] + 2.0, R, t[i]);
                        if (K >= 4) t[i] = fma(w[i] + 4.0, (R * R) * C2, t[i]);
                        if (i < 4) R *= eta;
                    }
                }
            }
#pragma unroll
            for (int i = 0; i < 5; ++i) {
                p[i] = A * t[i];
                if (i < 4) {
                    A *= ga;
                    ga *= de;
                }
            }
        } else {
            for (int k = -((K - 1) / 2); k <= K / 2; ++k) {
                const double two_k = (double)(2 * k);
#pragma unroll
                for (int i = 0; i < NW; ++i) {
                    const double u = w[i] + two_k;
                    p[i] = fma(u, exp_tab<true>(-(u * u) * arg, G.etab), p[i]);
                }
            }
        }
        // 1/sqrt(2 pi tt^3) = inv_tt * sqrt(inv_tt / (2 pi))
        post = inv_tt * (inv_tt * rsqrt_fast((2.0 * kPi) * inv_tt));
    } else {  // large-time, :59-63: terms k q^(k^2) sin(k pi w), ascending k
        const int K = -code;
        double sv[NW];
        sines(1, sv);
#pragma unroll
        for (int i = 0; i < NW; ++i) p[i] = q * sv[i];
        if (K >= 2) {
            const double q2 = q * q;
            double odd = q * q2, pw = q * odd;  // q^3, q^4
            double e = 2.0 * pw;
            sines(2, sv);
#pragma unroll
            for (int i = 0; i < NW; ++i) p[i] = fma(e, sv[i], p[i]);
            if (K >= 3) {
                odd *= q2; pw *= odd;  // q^5, q^9
                e = 3.0 * pw;
                sines(3, sv);
#pragma unroll
                for (int i = 0; i < NW; ++i) p[i] = fma(e, sv[i], p[i]);
                if (K >= 4) {
                    odd *= q2; pw *= odd;  // q^7, q^16
                    e = 4.0 * pw;
                    sines(4, sv);
#pragma unroll
                    for (int i = 0; i < NW; ++i) p[i] = fma(e, sv[i], p[i]);
                    if (KMAX > 4) {
                        for (int k = 5; k <= K; ++k) {  // K <= KMAX: the launch picked KMAX >= max K
                            odd *= q2; pw *= odd;
                            e = (double)k * pw;
                            sines(k, sv);
#pragma unroll
                            for (int i = 0; i < NW; ++i) p[i] = fma(e, sv[i], p[i]);
                        }
                    }
                }
            }
        }
        post = kPi;
    }
    // sv marginal, src/pdf.pxi:85,102: exp(((a sv w)^2 - 2 a v w - v^2 x) / (2 sv^2 x + 2)) / sqrt(sv^2 x + 1) / a^2.
    // The numerator is quadratic in the start-point index (equally spaced w_i), so the five exponentials are
    // E_0, E_0 g, E_0 g (g r2), ...: three exps and eight multiplications instead of five exps.
    double r = 1.0, d = 0.5;
    const bool sv_zero = G.sv == 0.0;
    if (!sv_zero) {
        r = rsqrt_fast(fma(G.sv_sq, xt, 1.0));
        d = 0.5 * (r * r);
    }
    const double scale = (post * r) * G.inv_a_sq_zt;
    // scale rides on the exponential chain: one multiplication instead of one per start point (|log| of
    // either factor is far below the exponent range: a^2 in [1e-100, 1e100], exponent in [-600, 300])
    double E = exp_tab<false>(fma(-c.v_sq, xt, c.c0) * d, G.etab) * scale;
    f[0] = p[0] * E;
    if (NW == 5) {
        double g = exp_tab<false>(c.dc0 * d, G.etab);
        const double r2 = exp_tab<false>(G.d2c[c.b] * d, G.etab);
#pragma unroll
        for (int i = 1; i < NW; ++i) {
            E *= g;
            if (i < NW - 1) g *= r2;
            f[i] = p[i] * E;
        }
    }
    if (!sv_zero) {  // exp(log(p) + ...) with p < 0 is NaN (src/pdf.pxi:102); never seen for err <= 1e-3
        int neg = __double2hiint(p[0]);
#pragma unroll
        for (int i = 1; i < NW; ++i) neg |= __double2hiint(p[i]);
        if (neg < 0) {
#pragma unroll
            for (int i = 0; i < NW; ++i)
                if (p[i] < 0.0) f[i] = __longlong_as_double(0x7ff8000000000000LL);
        }
    }
}

// The inner z-integration at decision time xt did not converge on its first level
// (src/integrate.pxi:99-112): walk deeper from the five values already known.  New start points are
// table nodes m/16 for the next two levels (sines from shared memory), arbitrary below that.
template <int KMAX>
__device__ __noinline__ double grid_refine_z(const GridSet<KMAX>& G, const GridTrial& c, double xt, int code,
                                             double q, const double f[5], double coarse, double tol,
                                             int levels, unsigned long long& evals)
{
    double post, qa;
    if (code > 0) {
        const double inv_tt = G.a_sq * rcp_fast(xt);
        qa = 0.5 * inv_tt;
        post = inv_tt * (inv_tt * rsqrt_fast((2.0 * kPi) * inv_tt));
    } else {
        qa = q;
        post = kPi;
    }
    double r = 1.0, d = 0.5;
    const bool sv_zero = G.sv == 0.0;
    if (!sv_zero) {
        r = rsqrt_fast(fma(G.sv_sq, xt, 1.0));
        d = 0.5 * (r * r);
    }
    const double scale = (post * r) * G.inv_a_sq_zt;
    const double vx = c.v_sq * xt;
    const double asv = G.a * G.sv, av2 = (2.0 * G.a) * c.v;
    const WTable<KMAX>& T = G.wt[c.b];
    unsigned long long extra = 0;
    auto g = [&](double w, int pos) {
        extra++;
        const int m = (pos & ((1 << kPosShift) - 1)) == 0 ? (pos >> kPosShift) : -1;
        double p = 0.0;
        if (code > 0) {
            const int K = code;
            for (int k = -((K - 1) / 2); k <= K / 2; ++k) {
                const double u = w + (double)(2 * k);
                p = fma(u, exp_tab<true>(-(u * u) * qa, G.etab), p);
            }
        } else {
            const int K = -code;
            const double q2 = qa * qa;
            double odd = qa, pw = qa;
            for (int k = 1; k <= K; ++k) {
                if (k > 1) {
                    odd *= q2;
                    pw *= odd;
                }
                const double kd = (double)k;
                const double sn = (m >= 0 && k <= KMAX) ? T.s[k - 1][m] : sin((kd * kPi) * w);
                p = fma(kd * pw, sn, p);
            }
        }
        const double t1 = asv * w;
        double val = (p * scale) * exp_tab<false>((fma(t1, t1, -av2 * w) - vx) * d, G.etab);
        if (!sv_zero && p < 0.0) val = __longlong_as_double(0x7ff8000000000000LL);
        return val;
    };
    Span stack[kMaxDepth];
    Span top;
    top.lo = T.w[0]; top.hi = T.w[16]; top.coarse = coarse;
    top.f_lo = f[0]; top.f_hi = f[4]; top.f_mid = f[2];
    top.tol = tol; top.levels = levels; top.ilo = 0; top.ihi = kPosWidth;
    const double res = simpson_walk(g, top, true, f[1], f[3], stack);
    evals += extra;
    return res;
}

// Trials outside the comfortable range: the v2 evaluator handles every corner (one out-of-line copy).
template <int KMAX>
__device__ __noinline__ double grid_fallback(double x, double v, const GridSet<KMAX>& G, const Numerics& nm,
                                             const SeriesTable* tab, unsigned long long& evals)
{
    return full_pdf_fast<KMAX>(x, v, G.raw[1], G.raw[2], G.raw[3], G.raw[4], G.raw[5], G.raw[6], nm, tab, G.wt,
                               evals);
}

// adaptiveSimpsons_1D over z at decision time xt (src/integrate.pxi:114-141), already / ZT -- or pdf_sv at
// the set's own start point when sz == 0.  xt is positive and inside the comfortable range.
// DEFER: the first-pass flavour -- a node whose inner integration needs refinement sets *defer instead of
// walking deeper (the caller re-evaluates the whole trial in its second pass), so that the node values never
// have their address taken and stay in registers.
template <int KMAX, bool DEFER = false>
__device__ __forceinline__ double grid_inner(const GridSet<KMAX>& G, const GridTrial& c, double xt, int code,
                                             double q, const Numerics& nm, const SeriesTable* tab,
                                             unsigned long long& extra, bool* defer = nullptr)
{
    if (G.sz != 0) {
        double f[5];
        grid_node<KMAX, 5>(G, c, xt, code, q, f);
        const double coarse = G.zw6 * ((f[0] + 4.0 * f[2]) + f[4]);
        const double left = G.zw12 * ((f[0] + 4.0 * f[1]) + f[2]);
        const double right = G.zw12 * ((f[2] + 4.0 * f[3]) + f[4]);
        const double fine = left + right;
        if (nm.n_sz <= 0 || fabs(fine - coarse) <= 15.0 * nm.simps_err) return fma(fine - coarse, 1.0 / 15.0, fine);
        if (DEFER) {
            *defer = true;
            return 0.0;
        }
        return grid_refine_z<KMAX>(G, c, xt, code, q, f, coarse, nm.simps_err, nm.n_sz, extra);
    }
    double f[1];
    grid_node<KMAX, 1>(G, c, xt, code, q, f);
    return f[0];
}

// Out-of-line copy for the paths outside the five-node loop (keeps the instruction footprint small).
template <int KMAX>
__device__ __noinline__ double grid_inner_ool(const GridSet<KMAX>& G, const GridTrial& c, double xt, int code,
                                              double q, const Numerics& nm, const SeriesTable* tab,
                                              unsigned long long& extra)
{
    return grid_inner<KMAX>(G, c, xt, code, q, nm, tab, extra);
}

// The outer t-integration did not converge on its first level (src/integrate.pxi:170-178): walk deeper
// with the same node evaluator.  *bad is set when a node leaves the comfortable range.
template <int KMAX>
__device__ __noinline__ double grid_refine_t(const GridSet<KMAX>& G, const GridTrial& c, double ax,
                                             const Numerics& nm, const SeriesTable* tab, const double F[5],
                                             double coarse, unsigned long long& evals, double xz, bool* bad,
                                             bool* tiny)
{
    const unsigned int per_node = G.sz != 0 ? 5 : 1;
    bool out_of_range = false, saw_tiny = false;
    auto g = [&](double tn, int) {
        const double xt = ax - tn;
        evals += per_node;
        if (!(xt > 0.0)) return 0.0;  // src/pdf.pxi:92-93
        if (xt < G.xt_lo[c.b]) {
            if (xt < xz) saw_tiny = true;
            else out_of_range = true;
            return 0.0;
        }
        const int code = grid_code(G, xt);
        const double q = code < 0 ? exp_tab<false>(G.qk * xt, G.etab) : 0.0;
        return grid_inner_ool<KMAX>(G, c, xt, code, q, nm, tab, evals) * G.inv_st;
    };
    Span top;
    top.lo = G.tn[0]; top.hi = G.tn[4]; top.coarse = coarse;
    top.f_lo = F[0]; top.f_hi = F[4]; top.f_mid = F[2];
    top.tol = nm.simps_err; top.levels = nm.n_st; top.ilo = 0; top.ihi = kPosWidth;
    Span stack[kMaxDepth];
    const double r = simpson_walk(g, top, true, F[1], F[3], stack);
    *bad = out_of_range;
    *tiny = saw_tiny;
    return r;
}

// full_pdf (src/pdf.pxi:104-146) for one trial of the CTA's set (v may differ per trial).  Returns through the v2 evaluator when the trial is outside the
// comfortable range (normalised time outside [2^-9, 128], extreme drift, NaN).
// DEFER (first pass of the likelihood kernel): trials that need anything beyond the first 5 x 5 level --
// inner or outer refinement, or the out-of-range fallback -- set *need and return 0 without counting
// evaluations; the kernel re-runs them, compacted, through full_pdf_grid_slow.
template <int KMAX, int STAGE = 0>
__device__ __forceinline__ double full_pdf_grid(double x, double v, const GridSet<KMAX>& G,
                                                const Numerics& nm, const SeriesTable* tab,
                                                unsigned long long& evals, bool* need = nullptr)
{
    constexpr bool DEFER = STAGE == 1;
    const double ax = fabs(x);
    if (!G.support || (ax - G.t_onset) < 0) return 0.0;  // src/pdf.pxi:111-113
    GridTrial c;
    c.b = x > 0 ? 1 : 0;
    c.v = c.b ? -v : v;
    c.v_sq = c.v * c.v;
    c.c0 = fma(-G.aw0[c.b], c.v, G.sw0[c.b]);
    c.dc0 = fma(-G.daw[c.b], c.v, G.dsw[c.b]);

    const double xt_max = ax - G.tn[0];
    // exponent of the sv marginal lies in [-(2 a |v| + v^2 x_t) / 2, (a sv w_4)^2 / 2 + a |v|]
    const double bnd = fma(c.v_sq, xt_max, (2.0 * G.a) * fabs(c.v));
    bool ok = G.ok && (xt_max <= G.xt_hi) && (bnd <= 600.0) && (nm.simps_err >= 0.0);
    // decision times in (0, xz): the node is negligible (see GridSet::xt_zero) and is taken as 0
    const double xz = (bnd <= 200.0 && nm.simps_err >= 1e-200) ? G.xt_zero[c.b] : 0.0;
    bool tiny = false;
    unsigned int n_eval = 0;
    unsigned long long extra = 0;
    double out = 0.0;
    if (ok) {
        const bool over_z = G.sz != 0;
        const double q0 = exp_tab<false>(G.qk * xt_max, G.etab);
        if (G.st == 0) {  // src/pdf.pxi:129,137-141: one decision time
            const double xt = ax - G.t;
            if (xt <= 0.0) {
                n_eval = over_z ? 5 : 1;  // src/pdf.pxi:92-93
            } else if (xt < G.xt_lo[c.b]) {
                ok = false;  // (a single negligible node would make the whole result negligible)
            } else {
                const int code = grid_code(G, xt);
                n_eval = over_z ? 5 : 1;
                out = grid_inner_ool<KMAX>(G, c, xt, code, q0, nm, tab, extra);
            }
        } else {  // :131-134, :143-146: five decision times, then the outer Simpson level
            double F[5];
            // First pass: a trial with a decision time at or before the onset, or below the comfortable range,
            // is left to the complete evaluator, so the node loop below carries no per-node range logic.
            if (DEFER && !(ax - G.tn[4] >= G.xt_lo[c.b])) {
                *need = true;
                return 0.0;
            }
            // the decision is monotone in x_t: if both ends agree every node does
            const int code_hi = grid_code(G, xt_max);
            int j_last = 4;
            if (!DEFER)
                while (j_last > 0 && !(ax - G.tn[j_last] > 0.0)) --j_last;
            const int code_lo = grid_code(G, ax - G.tn[j_last]);
            const bool uniform = code_hi == code_lo;
#pragma unroll(DEFER ? kUnrollJ : 1)
            for (int j = 0; j < 5; ++j) {
                const double xt = ax - G.tn[j];
                n_eval += over_z ? 5 : 1;
                double val = 0.0;
                if (DEFER) {
                    const int code = uniform ? code_hi : grid_code(G, xt);
                    const double q = q0 * G.rho[j];
                    if (!*need) val = grid_inner<KMAX, DEFER>(G, c, xt, code, q, nm, tab, extra, need);
                } else
                if (xt > 0.0) {
                    const bool live = !(xt < G.xt_lo[c.b]);
                    if (!live) {
                        if (xt < xz) tiny = true;
                        else ok = false;
                    }
                    const int code = uniform ? code_hi : grid_code(G, xt);
                    const double q = q0 * G.rho[j];
                    // (second pass: the out-of-line copy that grid_refine_t calls anyway -- one node evaluator
                    // in the instruction cache instead of two)
                    if (ok && live && !(DEFER && *need))
                        val = STAGE == 2 ? grid_inner_ool<KMAX>(G, c, xt, code, q, nm, tab, extra)
                                         : grid_inner<KMAX, DEFER>(G, c, xt, code, q, nm, tab, extra, need);
                }
                F[j] = val * G.inv_st;
            }
            if (DEFER && *need) return 0.0;
            if (ok) {
                const double coarse = G.tw6 * ((F[0] + 4.0 * F[2]) + F[4]);
                const double left = G.tw12 * ((F[0] + 4.0 * F[1]) + F[2]);
                const double right = G.tw12 * ((F[2] + 4.0 * F[3]) + F[4]);
                const double fine = left + right;
                if (nm.n_st <= 0 || fabs(fine - coarse) <= 15.0 * nm.simps_err)
                    out = fma(fine - coarse, 1.0 / 15.0, fine);
                else if (DEFER) {
                    *need = true;
                    return 0.0;
                } else {
                    bool bad = false, tiny2 = false;
                    out = grid_refine_t<KMAX>(G, c, ax, nm, tab, F, coarse, extra, xz, &bad, &tiny2);
                    ok = !bad;
                    tiny |= tiny2;
                }
                // Negligible nodes were taken as 0: fine unless the whole result is that small too -- and even
                // then when the caller's outlier term swamps it (fl(out (1 - p) + w p) = w p for any out <=
                // 1e-150 once w p > 1e-120; the recursion tree is the minimal one either way).
                if (tiny && !(fabs(out) > 1e-150) && !(fabs(out) <= 1e-150 && G.mix_floor > 1e-120)) ok = false;
            }
        }
    }
    if (!ok) {
        if (STAGE != 0) {  // stage 1: to the complete evaluator; stage 2: to the out-of-range fallback
            *need = true;
            return 0.0;
        }
        return grid_fallback<KMAX>(x, v, G, nm, tab, evals);
    }
    evals += n_eval + extra;
    return out;
}

// Second pass: the complete evaluator, one out-of-line copy.  A trial outside the comfortable range sets
// *fallback instead of walking through the v2 evaluator on its own while the rest of its warp idles: the
// kernel collects those and runs grid_fallback over them with full warps (third stage).
template <int KMAX>
__device__ __noinline__ double full_pdf_grid_slow(double x, double v, const GridSet<KMAX>& G, const Numerics& nm,
                                                  const SeriesTable* tab, unsigned long long& evals, bool* fallback)
{
    return full_pdf_grid<KMAX, 2>(x, v, G, nm, tab, evals, fallback);
}

}  // namespace hddm
