/*
 * cdfdif_oracle.c -- CPU restatement of the reference's full-DDM cumulative distribution.
 *
 * TEST INFRASTRUCTURE ONLY (see wfpt_oracle.c header): the parity checker for the CUDA
 * cdfdif kernel.  Never linked, imported or executed by anything under hddm_b200/.
 *
 * Restates src/cdfdif.c:59-221 (Tuerlinckx 2004 CDF: 6-point Gauss-Hermite over drift x
 * 6-point Gauss-Legendre over start point, Fourier series to 5000 terms with a 3-term
 * history stopping rule) and src/cdfdif_wrapper.pyx:16-52 (DMAT rescale, signed-axis CDF,
 * outlier mixture).  The reference's tests hold NO known-answer vector for this function
 * (SURVEY.md section 4), so parity is pinned by fixtures generated from the compiled
 * reference (the tests/golden npz fixtures, made by tests/golden/make_golden.py) and by live
 * bit-for-bit comparison against oracle/_ref when present.
 *
 * The algorithm is organised differently from the reference (quadrature rule set up once in
 * a struct, one helper per series) but every floating-point expression keeps the reference's
 * evaluation order so results are bit-identical under -O2 -ffp-contract=off.
 */
#include <math.h>
#include <stdint.h>

#define WO_PI 3.1415926535897932384626433832795028841971693993751
#define WO_NODES 6
#define WO_MAX_TERMS 5000
#define WO_TINY 1e-29    /* src/cdfdif.c:67 delta */
#define WO_NEARZERO 1e-7 /* src/cdfdif.c:68 epsilon */
#define WO_MIN_RT 0.001  /* src/cdfdif.c:69 */

/* Gauss-Hermite (drift) and Gauss-Legendre (start point) rules, src/cdfdif.c:78-81 */
static const double GH_X[WO_NODES] = {-2.3506049736744922818, -1.3358490740136970132,
                                      -.43607741192761650950, .43607741192761650950,
                                      1.3358490740136970132,  2.3506049736744922818};
static const double GH_W[WO_NODES] = {.45300099055088421593e-2, .15706732032114842368,
                                      .72462959522439207571,    .72462959522439207571,
                                      .15706732032114842368,    .45300099055088421593e-2};
static const double GL_X[WO_NODES] = {-.93246951420315193904, -.66120938646626381541,
                                      -.23861918608319693247, .23861918608319712676,
                                      .66120938646626459256,  .93246951420315160597};
static const double GL_W[WO_NODES] = {.17132449237917049545, .36076157304813916138,
                                      .46791393457269092604, .46791393457269092604,
                                      .36076157304813843973, .17132449237917132812};

typedef struct {
    double drift[WO_NODES], drift_w[WO_NODES]; /* scaled GH nodes/weights, :84-90 */
    double start[WO_NODES];                    /* scaled GL nodes, :91-92 */
} wo_rule;

typedef struct {
    double h0, h1, h2;
} wo_hist;

static long long g_series_terms = 0; /* instrumentation: total series iterations */
long long wo_cdf_series_terms(void) { return g_series_terms; }
void wo_cdf_series_terms_reset(void) { g_series_terms = 0; }

static int series_settled(const wo_hist *s)
{ /* src/cdfdif.c:144-145 (and :177-178, :202-203) */
    return (fabs(s->h0 - s->h1) < WO_TINY) && (fabs(s->h1 - s->h2) < WO_TINY) && (s->h2 > 0);
}

/* src/cdfdif.c:96-113: probability of the upper-coded boundary by 6x6 quadrature */
static double boundary_probability(const wo_rule *q, double a)
{
    double over_z = 0;
    int i, m;
    for (i = 0; i < WO_NODES; i++) {
        double over_nu = 0;
        for (m = 0; m < WO_NODES; m++) {
            if (fabs(q->drift[m]) > WO_NEARZERO)
                over_nu += (exp(-200 * q->start[i] * q->drift[m]) - 1) /
                           (exp(-200 * a * q->drift[m]) - 1) * q->drift_w[m];
            else
                over_nu += q->start[i] / a * q->drift_w[m];
        }
        over_z += over_nu * GL_W[i] / 2;
    }
    return over_z;
}

/* src/cdfdif.c:123-147: t beyond the whole Ter window; z and Ter integrated in closed form */
static double series_past_window(const wo_rule *q, int x, double a, double a2, double Z_U,
                                 double Z_L, double t, double upper_t, double lower_t)
{
    wo_hist s = {0, 0, 0};
    int v, m;
    for (v = 0; v < WO_MAX_TERMS; v++) {
        double over_nu = 0, sifa = WO_PI * v / a;
        s.h0 = s.h1;
        s.h1 = s.h2;
        g_series_terms++;
        for (m = 0; m < WO_NODES; m++) {
            double g = q->drift[m], wgt = q->drift_w[m];
            double denom = (100 * g * g + (WO_PI * WO_PI) * (v * v) / (100 * a2));
            double upp = exp((2 * x - 1) * Z_U * g * 100 - 3 * log(denom) + log(wgt) - 2 * log(100));
            double low = exp((2 * x - 1) * Z_L * g * 100 - 3 * log(denom) + log(wgt) - 2 * log(100));
            double fact = upp * ((2 * x - 1) * g * sin(sifa * Z_U) * 100 - sifa * cos(sifa * Z_U)) -
                          low * ((2 * x - 1) * g * sin(sifa * Z_L) * 100 - sifa * cos(sifa * Z_L));
            double exdif = exp((-.5 * denom * (t - upper_t)) +
                               log(1 - exp(-.5 * denom * (upper_t - lower_t))));
            over_nu += fact * exdif;
        }
        s.h2 = s.h1 + v * over_nu;
        if (series_settled(&s)) break;
    }
    return s.h2;
}

/* src/cdfdif.c:159-183: t inside the Ter window, drift node away from zero */
static double inside_window_drift_node(const wo_rule *q, int x, double g, double a, double a2,
                                       double st, double t, double lower_t)
{
    double over_z = 0;
    int i, v;
    for (i = 0; i < WO_NODES; i++) {
        double zzz = (a - q->start[i]) * x + q->start[i] * (1 - x);
        double ser = -((a * a2) / ((1 - 2 * x) * g * WO_PI * .01)) * sinh(zzz * (1 - 2 * x) * g / .01) /
                         (sinh((1 - 2 * x) * g * a / .01) * sinh((1 - 2 * x) * g * a / .01)) +
                     (zzz * a2) / ((1 - 2 * x) * g * WO_PI * .01) *
                         cosh((a - zzz) * (1 - 2 * x) * g / .01) / sinh((1 - 2 * x) * g * a / .01);
        wo_hist s = {0, 0, 0};
        for (v = 0; v < WO_MAX_TERMS; v++) {
            double sifa = WO_PI * v / a;
            double denom = (g * g * 100 + (WO_PI * v) * (WO_PI * v) / (a2 * 100));
            s.h0 = s.h1;
            s.h1 = s.h2;
            g_series_terms++;
            s.h2 = s.h1 + v * sin(sifa * zzz) * exp(-.5 * denom * (t - lower_t) - 2 * log(denom));
            if (series_settled(&s)) break;
        }
        over_z += .5 * GL_W[i] * (ser - 4 * s.h2) * (WO_PI / 100) / (a2 * st) *
                  exp((2 * x - 1) * zzz * g * 100);
    }
    return over_z;
}

/* src/cdfdif.c:185-207: t inside the Ter window, drift node numerically zero */
static double inside_window_zero_drift(double a, double a2, double Z_U, double Z_L, double sZ,
                                       double st, double t, double lower_t)
{
    wo_hist s = {0, 0, 0};
    double su = -(Z_U * Z_U) / (12 * a2) + (Z_U * Z_U * Z_U) / (12 * a * a2) -
                (Z_U * Z_U * Z_U * Z_U) / (48 * a2 * a2);
    double sl = -(Z_L * Z_L) / (12 * a2) + (Z_L * Z_L * Z_L) / (12 * a * a2) -
                (Z_L * Z_L * Z_L * Z_L) / (48 * a2 * a2);
    int v;
    for (v = 1; v < WO_MAX_TERMS; v++) {
        double sifa = WO_PI * v / a;
        double denom = (WO_PI * v) * (WO_PI * v) / (a2 * 100);
        s.h0 = s.h1;
        s.h1 = s.h2;
        g_series_terms++;
        s.h2 = s.h1 + 1 / (WO_PI * WO_PI * WO_PI * WO_PI * v * v * v * v) *
                          (cos(sifa * Z_L) - cos(sifa * Z_U)) * exp(-.5 * denom * (t - lower_t));
        if (series_settled(&s)) break;
    }
    return 400 * a2 * a * (sl - su - s.h2) / (st * sZ);
}

/* src/cdfdif.c:59-221  cdfdif(t, x, par[7], *prob) */
double wo_cdfdif(double t, int x, const double *par, double *prob)
{
    const double a = par[0], Ter = par[1], eta = par[2], z = par[3], sZ = par[4], st = par[5],
                 nu = par[6], a2 = a * a;
    const double Z_U = (1 - x) * z + x * (a - z) + sZ / 2;
    const double Z_L = (1 - x) * z + x * (a - z) - sZ / 2;
    const double lower_t = Ter - st / 2;
    double upper_t, p1, p0, F = 0.0;
    wo_rule q;
    int i, m;

    for (i = 0; i < WO_NODES; i++) {
        q.drift[i] = 1.41421356237309505 * GH_X[i] * eta + nu;
        q.drift_w[i] = GH_W[i] / 1.772453850905515882;
    }
    for (i = 0; i < WO_NODES; i++) q.start[i] = (.5 * sZ * GL_X[i]) + z;

    *prob = boundary_probability(&q, a);

    if (t - Ter + st / 2 > WO_MIN_RT) {
        upper_t = t < Ter + st / 2 ? t : Ter + st / 2;
        p1 = *prob * (upper_t - lower_t) / st;
        p0 = (1 - *prob) * (upper_t - lower_t) / st;
        if (t > Ter + st / 2) {
            double tail = series_past_window(&q, x, a, a2, Z_U, Z_L, t, upper_t, lower_t);
            F = (p0 * (1 - x) + p1 * x) - tail * 4 * WO_PI / (a2 * sZ * st);
        } else if (t <= Ter + st / 2) {
            double over_nu = 0;
            for (m = 0; m < WO_NODES; m++) {
                double part;
                if (fabs(q.drift[m]) > WO_NEARZERO)
                    part = inside_window_drift_node(&q, x, q.drift[m], a, a2, st, t, lower_t);
                else
                    part = inside_window_zero_drift(a, a2, Z_U, Z_L, sZ, st, t, lower_t);
                over_nu += part * q.drift_w[m];
            }
            F = (p0 * (1 - x) + p1 * x) - over_nu;
        }
    } else if (t - Ter + st / 2 <= WO_MIN_RT) {
        F = 0;
    }
    return F > WO_TINY ? F : 0;
}

/* src/cdfdif_wrapper.pyx:16-52  dmat_cdf_array.  Status: 0 ok, 1 = parameters out of support
 * (reference: ValueError :23-25), 2 = |x| >= 1/(2 w_outlier) with p_outlier > 0 (reference:
 * AssertionError :20-21).  w_outlier == 0 divides by zero in the mixture (:12) -- the Python
 * callers guard it; here it follows IEEE. */
int wo_dmat_cdf_array(const double *x, int64_t n, double v, double sv, double a, double z,
                      double sz, double t, double st, double p_outlier, double w_outlier, double *y)
{
    double par[7], p_up;
    int64_t i;
    if (p_outlier > 0) {
        double peak = 0;
        for (i = 0; i < n; i++)
            if (fabs(x[i]) > peak) peak = fabs(x[i]);
        if (n > 0 && !(peak < (1. / (2 * w_outlier)))) return 2;
    }
    if ((sv < 0) || (a <= 0) || (z < 0) || (z > 1) || (sz < 0) || (sz > 1) || (z + sz / 2. > 1) ||
        (z - sz / 2. < 0) || (t - st / 2. < 0) || (t < 0) || (st < 0) ||
        !((p_outlier >= 0) & (p_outlier <= 1)))
        return 1;
    par[0] = a / 10.;
    par[1] = t;
    par[2] = sv / 10. + 1e-10;
    par[3] = z * (a / 10.);
    par[4] = sz * (a / 10.) + 1e-10;
    par[5] = st + 1e-10;
    par[6] = v / 10.;
    for (i = 0; i < n; i++) {
        int upper = (int)(x[i] > 0);
        double sgn = (x[i] > 0) - (x[i] < 0);
        double F = wo_cdfdif(fabs(x[i]), upper, par, &p_up);
        F = (1 - p_up) + sgn * F;
        y[i] = F * (1 - p_outlier) + (x[i] + (1. / (2 * w_outlier))) * w_outlier * p_outlier;
    }
    return 0;
}
