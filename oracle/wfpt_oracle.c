/*
 * wfpt_oracle.c -- CPU restatement of the reference's Wiener first-passage-time hot path.
 *
 * TEST INFRASTRUCTURE ONLY.  This file is the parity checker for the CUDA path: only tests/,
 * __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load it.
 * Nothing under hddm_b200/ links, imports or executes it.
 *
 * Parity status: PINNED.  tests/test_oracle.py checks this restatement against
 *   (1) the reference's 20 MATLAB golden vectors (hddm/tests/matlab_values.py:3-162, via
 *       tests/golden/matlab_wfpt.json),
 *   (2) fixtures generated from the compiled, unmodified reference (oracle/_ref, built by
 *       oracle/build_ref.py from /root/reference/src) committed under tests/golden/, and
 *   (3) live bit-for-bit comparison against oracle/_ref whenever that build is present.
 *
 * Every function cites the reference file:line (relative to /root/reference) it restates.
 * The floating-point operation ORDER follows the C that Cython generates for the reference
 * (pow(x,2.0) for **2, pow(tt,3.0), left-to-right products) so results are bit-identical
 * to the stock -O2 reference build when compiled with -O2 -ffp-contract=off.
 *
 * Plain C99 + libm.  All state is per-call except an instrumentation counter of pdf_sv
 * evaluations (the unit SURVEY.md section 8(d) uses for algorithmic FLOP accounting).
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>

#ifndef M_PI
#define M_PI 3.14159265358979323846
#endif

static __thread long long g_pdfsv_calls = 0;

long long wo_pdfsv_calls(void) { return g_pdfsv_calls; }
void wo_pdfsv_calls_reset(void) { g_pdfsv_calls = 0; }

typedef struct {
    double v, sv, a, err;
} wo_core;

static double dmax(double p, double q) { return p < q ? q : p; } /* std::max<double>: (a<b)?b:a */

/* src/pdf.pxi:28-65  ftt_01w -- Navarro & Fuss (2009) f(tt | 0, 1, w) */
double wo_ftt_01w(double tt, double w, double err)
{
    double kl, ks, dens = 0.0;
    int k, nterms;

    /* terms needed by the large-time expansion, src/pdf.pxi:36-40 */
    if (M_PI * tt * err < 1.0) {
        kl = sqrt(-2.0 * log(M_PI * tt * err) / (pow(M_PI, 2.0) * tt));
        kl = dmax(kl, 1. / (M_PI * sqrt(tt)));
    } else {
        kl = 1. / (M_PI * sqrt(tt));
    }
    /* terms needed by the small-time expansion, src/pdf.pxi:43-47 */
    if (2.0 * sqrt(2.0 * M_PI * tt) * err < 1.0) {
        ks = 2.0 + sqrt(-2.0 * tt * log(2.0 * sqrt(2.0 * M_PI * tt) * err));
        ks = dmax(ks, sqrt(tt) + 1.0);
    } else {
        ks = 2.0;
    }

    if (ks < kl) { /* small-time series, src/pdf.pxi:51-57 */
        int first, last;
        nterms = (int)ceil(ks);
        first = (int)(-floor((nterms - 1) / 2.));
        last = (int)ceil((nterms - 1) / 2.);
        for (k = first; k <= last; k++) {
            double u = w + (2 * k);
            dens = dens + u * exp(((-pow(u, 2.0)) / 2.0) / tt);
        }
        dens = dens / sqrt((2.0 * M_PI) * pow(tt, 3.0));
    } else { /* large-time series, src/pdf.pxi:59-63 */
        nterms = (int)ceil(kl);
        for (k = 1; k <= nterms; k++) {
            dens = dens + (k * exp((((-pow((double)k, 2.0)) * pow(M_PI, 2.0)) * tt) / 2.0)) *
                              sin((k * M_PI) * w);
        }
        dens = dens * M_PI;
    }
    return dens;
}

/* src/pdf.pxi:36-54,60  the (branch, K) decision of ftt_01w alone: +K small-time series with K
 * terms, -K large-time series with K terms (0: large-time, empty sum).  Same expressions as
 * wo_ftt_01w above; exists so tests can check the CUDA path's host-built selection table. */
int wo_series_code(double tt, double err)
{
    double kl, ks;
    if (M_PI * tt * err < 1.0) {
        kl = sqrt(-2.0 * log(M_PI * tt * err) / (pow(M_PI, 2.0) * tt));
        kl = dmax(kl, 1. / (M_PI * sqrt(tt)));
    } else {
        kl = 1. / (M_PI * sqrt(tt));
    }
    if (2.0 * sqrt(2.0 * M_PI * tt) * err < 1.0) {
        ks = 2.0 + sqrt(-2.0 * tt * log(2.0 * sqrt(2.0 * M_PI * tt) * err));
        ks = dmax(ks, sqrt(tt) + 1.0);
    } else {
        ks = 2.0;
    }
    if (ks < kl) return (int)ceil(ks);
    return -(int)ceil(kl);
}

/* src/pdf.pxi:67-72  prob_ub -- probability of absorbing at the upper boundary */
double wo_prob_ub(double v, double a, double z)
{
    if (v == 0.0) return z;
    return (exp(((-2.0 * a) * z) * v) - 1.0) / (exp((-2.0 * a) * v) - 1.0);
}

/* src/pdf.pxi:74-85  pdf -- f(x | v, a, w) without drift variability */
static double density_fixed_drift(double x, double v, double a, double w, double err)
{
    double tt, p;
    if (x <= 0.0) return 0.0;
    tt = x / pow(a, 2.0);
    p = wo_ftt_01w(tt, w, err);
    return (p * exp(((-v) * a) * w - (pow(v, 2.0) * x) / 2.)) / pow(a, 2.0);
}

/* src/pdf.pxi:87-102  pdf_sv -- analytic marginal over normally distributed drift */
double wo_pdf_sv(double x, double v, double sv, double a, double z, double err)
{
    double tt, p;
    g_pdfsv_calls++;
    if (x <= 0.0) return 0.0;
    if (sv == 0.0) return density_fixed_drift(x, v, a, z, err);
    tt = x / pow(a, 2.0);
    p = wo_ftt_01w(tt, z, err);
    return (exp(log(p) + ((pow((a * z) * sv, 2.0) - ((2.0 * a) * v) * z) - pow(v, 2.0) * x) /
                             ((2.0 * pow(sv, 2.0)) * x + 2.0)) /
            sqrt(pow(sv, 2.0) * x + 1.0)) /
           pow(a, 2.0);
}

/* src/integrate.pxi:12-45  simpson_1D -- fixed composite Simpson along z (n_st==0) or t */
static double fixed_simpson_line(double x, const wo_core *c, double z, double t, double lb_z,
                                 double ub_z, int n_sz, double lb_t, double ub_t, int n_st)
{
    double ht, hz, acc, y = 0.0;
    int i, n = n_st > n_sz ? n_st : n_sz;
    if (n_st == 0) {
        hz = (ub_z - lb_z) / n;
        ht = 0.0;
        lb_t = t;
        ub_t = t;
    } else {
        hz = 0.0;
        ht = (ub_t - lb_t) / n;
        lb_z = z;
        ub_z = z;
    }
    acc = wo_pdf_sv(x - lb_t, c->v, c->sv, c->a, lb_z, c->err);
    for (i = 1; i <= n; i++) {
        double zi = lb_z + hz * i;
        double ti = lb_t + ht * i;
        y = wo_pdf_sv(x - ti, c->v, c->sv, c->a, zi, c->err);
        acc = (i & 1) ? acc + 4.0 * y : acc + 2.0 * y;
    }
    acc = acc - y; /* the last node was weighted 2, should be 1 (src/integrate.pxi:42) */
    acc = acc / ((ub_t - lb_t) + (ub_z - lb_z));
    return ((ht + hz) * acc) / 3.0;
}

/* src/integrate.pxi:47-70  simpson_2D -- Simpson over t of Simpson over z */
static double fixed_simpson_plane(double x, const wo_core *c, double z, double lb_z, double ub_z,
                                  int n_sz, double lb_t, double ub_t, int n_st)
{
    double ht = (ub_t - lb_t) / n_st, acc, y = 0.0;
    int j;
    acc = fixed_simpson_line(x, c, z, lb_t, lb_z, ub_z, n_sz, 0, 0, 0);
    for (j = 1; j <= n_st; j++) {
        double tj = lb_t + ht * j;
        y = fixed_simpson_line(x, c, z, tj, lb_z, ub_z, n_sz, 0, 0, 0);
        acc = (j & 1) ? acc + 4.0 * y : acc + 2.0 * y;
    }
    acc = acc - y;
    acc = acc / (ub_t - lb_t);
    return (ht * acc) / 3.0;
}

/* src/integrate.pxi:72-112  adaptiveSimpsonsAux -- one refinement step, left then right */
static double refine_line(double x, const wo_core *c, double z, double t, double lb_z, double ub_z,
                          double lb_t, double ub_t, double width, double tol, double coarse,
                          double f_lo, double f_hi, double f_mid, int levels_left)
{
    double z_c, z_d, z_e, t_c, t_d, t_e, h, fd, fe, left, right, fine;
    if ((ub_t - lb_t) == 0) { /* integrate over z */
        h = ub_z - lb_z;
        z_c = (ub_z + lb_z) / 2.;
        z_d = (lb_z + z_c) / 2.;
        z_e = (z_c + ub_z) / 2.;
        t_c = t_d = t_e = t;
    } else { /* integrate over t */
        h = ub_t - lb_t;
        t_c = (ub_t + lb_t) / 2.;
        t_d = (lb_t + t_c) / 2.;
        t_e = (t_c + ub_t) / 2.;
        z_c = z_d = z_e = z;
    }
    fd = wo_pdf_sv(x - t_d, c->v, c->sv, c->a, z_d, c->err) / width;
    fe = wo_pdf_sv(x - t_e, c->v, c->sv, c->a, z_e, c->err) / width;
    left = (h / 12.0) * ((f_lo + 4.0 * fd) + f_mid);
    right = (h / 12.0) * ((f_mid + 4.0 * fe) + f_hi);
    fine = left + right;
    if (levels_left <= 0 || fabs(fine - coarse) <= 15.0 * tol)
        return fine + (fine - coarse) / 15.0;
    return refine_line(x, c, z, t, lb_z, z_c, lb_t, t_c, width, tol / 2, left, f_lo, f_mid, fd,
                       levels_left - 1) +
           refine_line(x, c, z, t, z_c, ub_z, t_c, ub_t, width, tol / 2, right, f_mid, f_hi, fe,
                       levels_left - 1);
}

/* src/integrate.pxi:114-141  adaptiveSimpsons_1D */
static double adaptive_line(double x, const wo_core *c, double z, double t, double lb_z,
                            double ub_z, double lb_t, double ub_t, double tol, int depth)
{
    double h, width, c_t, c_z, f_lo, f_hi, f_mid, coarse;
    if ((ub_t - lb_t) == 0) {
        lb_t = t;
        ub_t = t;
        h = ub_z - lb_z;
    } else {
        h = ub_t - lb_t;
        lb_z = z;
        ub_z = z;
    }
    width = h;
    c_t = (lb_t + ub_t) / 2.;
    c_z = (lb_z + ub_z) / 2.;
    f_lo = wo_pdf_sv(x - lb_t, c->v, c->sv, c->a, lb_z, c->err) / width;
    f_hi = wo_pdf_sv(x - ub_t, c->v, c->sv, c->a, ub_z, c->err) / width;
    f_mid = wo_pdf_sv(x - c_t, c->v, c->sv, c->a, c_z, c->err) / width;
    coarse = (h / 6.0) * ((f_lo + 4.0 * f_mid) + f_hi);
    return refine_line(x, c, z, t, lb_z, ub_z, lb_t, ub_t, width, tol, coarse, f_lo, f_hi, f_mid,
                       depth);
}

/* src/integrate.pxi:143-178  adaptiveSimpsonsAux_2D -- outer refinement over t, inner over z */
static double refine_plane(double x, const wo_core *c, double z, double tol_z, double lb_z,
                           double ub_z, double lb_t, double ub_t, double st, double tol_t,
                           double coarse, double f_lo, double f_hi, double f_mid, int depth_z,
                           int levels_left)
{
    double t_c = (ub_t + lb_t) / 2.;
    double t_d = (lb_t + t_c) / 2.;
    double t_e = (t_c + ub_t) / 2.;
    double h = ub_t - lb_t;
    double fd = adaptive_line(x, c, z, t_d, lb_z, ub_z, 0, 0, tol_z, depth_z) / st;
    double fe = adaptive_line(x, c, z, t_e, lb_z, ub_z, 0, 0, tol_z, depth_z) / st;
    double left = (h / 12.0) * ((f_lo + 4.0 * fd) + f_mid);
    double right = (h / 12.0) * ((f_mid + 4.0 * fe) + f_hi);
    double fine = left + right;
    if (levels_left <= 0 || fabs(fine - coarse) <= 15.0 * tol_t)
        return fine + (fine - coarse) / 15.0;
    return refine_plane(x, c, z, tol_z, lb_z, ub_z, lb_t, t_c, st, tol_t / 2, left, f_lo, f_mid,
                        fd, depth_z, levels_left - 1) +
           refine_plane(x, c, z, tol_z, lb_z, ub_z, t_c, ub_t, st, tol_t / 2, right, f_mid, f_hi,
                        fe, depth_z, levels_left - 1);
}

/* src/integrate.pxi:181-206  adaptiveSimpsons_2D */
static double adaptive_plane(double x, const wo_core *c, double z, double lb_z, double ub_z,
                             double lb_t, double ub_t, double tol, int depth_z, int depth_t)
{
    double h = ub_t - lb_t, st = ub_t - lb_t;
    double f_lo = adaptive_line(x, c, z, lb_t, lb_z, ub_z, 0, 0, tol, depth_z) / st;
    double f_hi = adaptive_line(x, c, z, ub_t, lb_z, ub_z, 0, 0, tol, depth_z) / st;
    double f_mid = adaptive_line(x, c, z, (lb_t + ub_t) / 2, lb_z, ub_z, 0, 0, tol, depth_z) / st;
    double coarse = (h / 6.0) * ((f_lo + 4.0 * f_mid) + f_hi);
    return refine_plane(x, c, z, tol, lb_z, ub_z, lb_t, ub_t, st, tol, coarse, f_lo, f_hi, f_mid,
                        depth_z, depth_t);
}

/* src/pdf.pxi:104-146  full_pdf -- support test, boundary flip, 4-way dispatch */
double wo_full_pdf(double x, double v, double sv, double a, double z, double sz, double t,
                   double st, double err, int n_st, int n_sz, int use_adaptive, double simps_err)
{
    wo_core c;
    if ((z < 0) || (z > 1) || (a < 0) || (t < 0) || (st < 0) || (sv < 0) || (sz < 0) || (sz > 1) ||
        ((fabs(x) - (t - st / 2.)) < 0) || (z + sz / 2. > 1) || (z - sz / 2. < 0) ||
        (t - st / 2. < 0))
        return 0.0;
    if (x > 0) {
        v = -v;
        z = 1. - z;
    }
    x = fabs(x);
    if (st < 1e-3) st = 0;
    if (sz < 1e-3) sz = 0;
    c.v = v; c.sv = sv; c.a = a; c.err = err;

    if (sz == 0) {
        if (st == 0) return wo_pdf_sv(x - t, v, sv, a, z, err);
        if (use_adaptive > 0)
            return adaptive_line(x, &c, z, t, z, z, t - st / 2., t + st / 2., simps_err, n_st);
        return fixed_simpson_line(x, &c, z, t, z, z, 0, t - st / 2., t + st / 2., n_st);
    }
    if (st == 0) {
        if (use_adaptive)
            return adaptive_line(x, &c, z, t, z - sz / 2., z + sz / 2., t, t, simps_err, n_sz);
        return fixed_simpson_line(x, &c, z, t, z - sz / 2., z + sz / 2., n_sz, t, t, 0);
    }
    if (use_adaptive)
        return adaptive_plane(x, &c, z, z - sz / 2., z + sz / 2., t - st / 2., t + st / 2.,
                              simps_err, n_sz, n_st);
    return fixed_simpson_plane(x, &c, z, z - sz / 2., z + sz / 2., n_sz, t - st / 2., t + st / 2.,
                               n_st);
}

/* src/wfpt.pyx:32-48  pdf_array.  Note: the reference takes log with numpy's vectorised
 * np.log, which may differ from libm log in the last ulp on AVX-512 hosts. */
void wo_pdf_array(const double *x, int64_t n, double v, double sv, double a, double z, double sz,
                  double t, double st, double err, int logp, int n_st, int n_sz, int use_adaptive,
                  double simps_err, double p_outlier, double w_outlier, double *y)
{
    int64_t i;
    for (i = 0; i < n; i++) {
        double d = wo_full_pdf(x[i], v, sv, a, z, sz, t, st, err, n_st, n_sz, use_adaptive,
                               simps_err);
        d = d * (1 - p_outlier) + (w_outlier * p_outlier);
        y[i] = logp == 1 ? log(d) : d;
    }
}

/* src/wfpt.pyx:54-76  wiener_like */
double wo_wiener_like(const double *x, int64_t n, double v, double sv, double a, double z,
                      double sz, double t, double st, double err, int n_st, int n_sz,
                      int use_adaptive, double simps_err, double p_outlier, double w_outlier)
{
    double total = 0.0, mix = w_outlier * p_outlier;
    int64_t i;
    if (!((p_outlier >= 0) & (p_outlier <= 1))) return -INFINITY;
    for (i = 0; i < n; i++) {
        double d = wo_full_pdf(x[i], v, sv, a, z, sz, t, st, err, n_st, n_sz, use_adaptive,
                               simps_err);
        d = d * (1 - p_outlier) + mix;
        if (d == 0) return -INFINITY;
        total += log(d);
    }
    return total;
}

/* src/wfpt.pyx:244-274  wiener_like_multi.  Each of the seven parameters is a (pointer, stride)
 * pair; stride 0 broadcasts a scalar, stride 1 reads a per-trial array (the `multi` names). */
double wo_wiener_like_multi(const double *x, int64_t n, const double *v, int64_t v_s,
                            const double *sv, int64_t sv_s, const double *a, int64_t a_s,
                            const double *z, int64_t z_s, const double *sz, int64_t sz_s,
                            const double *t, int64_t t_s, const double *st, int64_t st_s,
                            double err, int n_st, int n_sz, int use_adaptive, double simps_err,
                            double p_outlier, double w_outlier)
{
    double total = 0.0, mix = w_outlier * p_outlier, d;
    int64_t i;
    for (i = 0; i < n; i++) {
        double vi = v[i * v_s], ai = a[i * a_s], zi = z[i * z_s];
        if (fabs(x[i]) != 999.) {
            d = wo_full_pdf(x[i], vi, sv[i * sv_s], ai, zi, sz[i * sz_s], t[i * t_s],
                            st[i * st_s], err, n_st, n_sz, use_adaptive, simps_err);
            d = d * (1 - p_outlier) + mix;
        } else if (x[i] == 999.) {
            d = wo_prob_ub(vi, ai, zi);
        } else {
            d = 1 - wo_prob_ub(vi, ai, zi);
        }
        total += log(d);
    }
    return total;
}

/* src/wfpt.pyx:357-386  wiener_like_contaminant */
double wo_wiener_like_contaminant(const double *x, const int32_t *cont_x, int64_t n, double v,
                                  double sv, double a, double z, double sz, double t, double st,
                                  double t_min, double t_max, double err, int n_st, int n_sz,
                                  int use_adaptive, double simps_err)
{
    double total = 0.0;
    int64_t i;
    int n_cont = 0;
    for (i = 0; i < n; i++) n_cont += cont_x[i];
    for (i = 0; i < n; i++) {
        if (cont_x[i] == 0) {
            double d = wo_full_pdf(x[i], v, sv, a, z, sz, t, st, err, n_st, n_sz, use_adaptive,
                                   simps_err);
            if (d == 0) return -INFINITY;
            total += log(d);
        }
    }
    total += n_cont * log(0.5 * 1. / (t_max - t_min));
    return total;
}

/* NEW entry point mirrored by the CUDA path (no reference counterpart): P parameter sets, each
 * evaluated with wiener_like semantics (src/wfpt.pyx:54-76) against its own trial segment
 * x[seg[p] .. seg[p+1]) (seg == NULL: every set sees all n trials).  params is SoA
 * [9][P]: v, sv, a, z, sz, t, st, p_outlier, w_outlier. */
void wo_wiener_like_batch(const double *x, int64_t n, const int64_t *seg, const double *params,
                          int64_t P, double err, int n_st, int n_sz, int use_adaptive,
                          double simps_err, double *logp)
{
    int64_t p;
    for (p = 0; p < P; p++) {
        int64_t lo = seg ? seg[p] : 0, hi = seg ? seg[p + 1] : n;
        logp[p] = wo_wiener_like(x + lo, hi - lo, params[0 * P + p], params[1 * P + p],
                                 params[2 * P + p], params[3 * P + p], params[4 * P + p],
                                 params[5 * P + p], params[6 * P + p], err, n_st, n_sz,
                                 use_adaptive, simps_err, params[7 * P + p], params[8 * P + p]);
    }
}

/* src/wfpt.pyx:389-414  gen_cdf_using_pdf, given the caller's grid x = linspace(-time,time,2N+1).
 * Returns 0, or 1 when the parameters are out of support (reference raises ValueError :395-398).
 * cumtrapz(y) with unit spacing = cumsum(1.0*(y[1:]+y[:-1])/2.0); cdf[0] keeps the raw pdf. */
int wo_gen_cdf_using_pdf(const double *xgrid, int64_t m, double v, double sv, double a, double z,
                         double sz, double t, double st, double err, int n_st, int n_sz,
                         int use_adaptive, double simps_err, double p_outlier, double w_outlier,
                         double *cdf)
{
    int64_t i;
    double run = 0.0, prev, last;
    if ((sv < 0) || (a <= 0) || (z < 0) || (z > 1) || (sz < 0) || (sz > 1) || (z + sz / 2. > 1) ||
        (z - sz / 2. < 0) || (t - st / 2. < 0) || (t < 0) || (st < 0) ||
        !((p_outlier >= 0) & (p_outlier <= 1)))
        return 1;
    wo_pdf_array(xgrid, m, v, sv, a, z, sz, t, st, err, 0, n_st, n_sz, use_adaptive, simps_err,
                 p_outlier, w_outlier, cdf);
    prev = cdf[0];
    for (i = 1; i < m; i++) {
        double cur = cdf[i];
        run += (1.0 * (cur + prev)) / 2.0;
        prev = cur;
        cdf[i] = run;
    }
    last = cdf[m - 1];
    for (i = 0; i < m; i++) cdf[i] /= last;
    return 0;
}

/* src/wfpt.pyx:323-354  gen_rts_from_cdf, with the caller's grid x = arange(lb,ub,dt) and the
 * caller's uniforms f (and u_delay when st != 0) in place of numpy's global RNG.
 * l_cdf (length m) is scratch/output: the normalised rectangle-rule table. */
void wo_gen_rts_from_cdf(const double *xgrid, int64_t m, double v, double sv, double a, double z,
                         double sz, double t, double st, const double *f, const double *u_delay,
                         int64_t samples, double *l_cdf, double *rts)
{
    int64_t i;
    double last;
    l_cdf[0] = 0;
    for (i = 1; i < m; i++)
        l_cdf[i] = l_cdf[i - 1] + wo_full_pdf(xgrid[i], v, sv, a, z, sz, 0, 0, 1e-4, 2, 2, 1, 1e-3);
    last = l_cdf[m - 1];
    for (i = 0; i < m; i++) l_cdf[i] /= last;
    for (i = 0; i < samples; i++) {
        /* np.searchsorted(l_cdf, f, side='left'): first index with l_cdf[idx] >= f */
        int64_t lo = 0, hi = m;
        double rt, sgn, shift;
        while (lo < hi) {
            int64_t mid = lo + (hi - lo) / 2;
            if (l_cdf[mid] < f[i]) lo = mid + 1; else hi = mid;
        }
        rt = xgrid[lo];
        sgn = (rt > 0) - (rt < 0);
        shift = (st == 0) ? t : (u_delay[i] * st + (t - st / 2.));
        rts[i] = rt + sgn * shift;
    }
}

/* ---- reinforcement-learning likelihoods -------------------------------- */

/* learning rate on the logit scale, src/wfpt.pyx:123-126: 2.718281828459**a / (1 + 2.718281828459**a)
 * (the reference's truncated e, lowered by Cython to pow()). */
static double rl_rate(double logit) { return pow(2.718281828459, logit) / (1 + pow(2.718281828459, logit)); }

/* Conditions in the order of np.unique(split_by) (ascending), each visited with its trials in the
 * caller's order (boolean-mask selection, src/wfpt.pyx:112-118).  next_condition returns 0 when
 * none is left; *cur is the condition just handled (start with have_cur = 0). */
static int next_condition(const int64_t *split_by, int64_t n, int have_cur, int64_t cur, int64_t *out)
{
    int found = 0;
    int64_t i, best = 0;
    for (i = 0; i < n; i++)
        if ((!have_cur || split_by[i] > cur) && (!found || split_by[i] < best)) {
            best = split_by[i];
            found = 1;
        }
    *out = best;
    return found;
}

/* src/wfpt.pyx:79-155  wiener_like_rlddm.  response is 0 (lower) / 1 (upper), int64 like the
 * reference's np.ndarray[long]. */
double wo_wiener_like_rlddm(const double *x, const int64_t *response, const double *feedback,
                            const int64_t *split_by, int64_t n, double q, double alpha,
                            double pos_alpha, double v, double sv, double a, double z, double sz,
                            double t, double st, double err, int n_st, int n_sz, int use_adaptive,
                            double simps_err, double p_outlier, double w_outlier)
{
    double total = 0.0, mix = w_outlier * p_outlier, qs[2], alfa, pos_alfa, d;
    int64_t s = 0, i;
    int have = 0;
    if (!((p_outlier >= 0) & (p_outlier <= 1))) return -INFINITY;
    pos_alfa = (pos_alpha == 100.00) ? alpha : pos_alpha; /* :106-109 */
    while (next_condition(split_by, n, have, s, &s)) {
        int first = 1;
        have = 1;
        qs[0] = q;
        qs[1] = q;
        for (i = 0; i < n; i++) {
            if (split_by[i] != s) continue;
            if (!first) { /* :134-140: no density for the first trial of a condition */
                d = wo_full_pdf(x[i], (qs[1] - qs[0]) * v, sv, a, z, sz, t, st, err, n_st, n_sz,
                                use_adaptive, simps_err);
                d = d * (1 - p_outlier) + mix;
                if (d == 0) return -INFINITY;
                total += log(d);
            }
            first = 0;
            /* :122-131, :145-154 */
            alfa = (feedback[i] > qs[response[i]]) ? rl_rate(pos_alfa) : rl_rate(alpha);
            qs[response[i]] = qs[response[i]] + alfa * (feedback[i] - qs[response[i]]);
        }
    }
    return total;
}

/* src/wfpt.pyx:158-241  wiener_like_rl: choice-only likelihood (softmax-like P(upper) of a unit
 * boundary separation process), same Q-learning updates. */
double wo_wiener_like_rl(const int64_t *response, const double *feedback, const int64_t *split_by,
                         int64_t n, double q, double alpha, double pos_alpha, double v, double z,
                         double p_outlier, double w_outlier)
{
    double total = 0.0, mix = w_outlier * p_outlier, qs[2], alfa, pos_alfa, d, drift;
    int64_t s = 0, i;
    int have = 0;
    if (!((p_outlier >= 0) & (p_outlier <= 1))) return -INFINITY;
    pos_alfa = (pos_alpha == 100.00) ? alpha : pos_alpha;
    while (next_condition(split_by, n, have, s, &s)) {
        int first = 1;
        have = 1;
        qs[0] = q;
        qs[1] = q;
        for (i = 0; i < n; i++) {
            if (split_by[i] != s) continue;
            if (!first) {
                drift = (qs[1] - qs[0]) * v;
                if (drift == 0) {
                    d = 0.5;
                } else if (response[i] == 1) { /* :213-216 */
                    d = (pow(2.718281828459, -2 * z * drift) - 1) / (pow(2.718281828459, -2 * drift) - 1);
                } else {
                    d = 1 - (pow(2.718281828459, -2 * z * drift) - 1) / (pow(2.718281828459, -2 * drift) - 1);
                }
                d = d * (1 - p_outlier) + mix;
                if (d == 0) return -INFINITY;
                total += log(d);
            }
            first = 0;
            alfa = (feedback[i] > qs[response[i]]) ? rl_rate(pos_alfa) : rl_rate(alpha);
            qs[response[i]] = qs[response[i]] + alfa * (feedback[i] - qs[response[i]]);
        }
    }
    return total;
}

/* src/wfpt.pyx:277-320  wiener_like_multi_rlddm: trials in the caller's order, Q values reset
 * where split_by changes between neighbours (:300-303), every trial scored (no first-trial skip),
 * single learning rate (no pos_alpha), no zero-density guard.  par order: v, sv, a, z, sz, t, st,
 * alpha as (pointer, stride) pairs (stride 0 = scalar, 1 = per trial: the `multi` names). */
double wo_wiener_like_multi_rlddm(const double *x, const int64_t *response, const double *feedback,
                                  const int64_t *split_by, int64_t n, double q,
                                  const double *const par[8], const int64_t stride[8], double err,
                                  int n_st, int n_sz, int use_adaptive, double simps_err,
                                  double p_outlier, double w_outlier)
{
    double total = 0.0, mix = w_outlier * p_outlier, qs[2], d, alfa;
    int64_t i;
    qs[0] = q;
    qs[1] = q;
    for (i = 0; i < n; i++) {
        if (i != 0 && split_by[i] != split_by[i - 1]) {
            qs[0] = q;
            qs[1] = q;
        }
        d = wo_full_pdf(x[i], par[0][i * stride[0]] * (qs[1] - qs[0]), par[1][i * stride[1]],
                        par[2][i * stride[2]], par[3][i * stride[3]], par[4][i * stride[4]],
                        par[5][i * stride[5]], par[6][i * stride[6]], err, n_st, n_sz, use_adaptive,
                        simps_err);
        d = d * (1 - p_outlier) + mix;
        total += log(d);
        alfa = rl_rate(par[7][i * stride[7]]);
        qs[response[i]] = qs[response[i]] + alfa * (feedback[i] - qs[response[i]]);
    }
    return total;
}
