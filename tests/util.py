"""Shared parity helpers.  Tolerances are the north_star's: 1e-10 relative per-trial (log-)density,
1e-9 relative summed log-likelihood, FP64; 0 / +-inf / NaN must match as exact classes."""
import numpy as np

RTOL_TRIAL = 1e-10
RTOL_SUM = 1e-9
ATOL_LOG = 1e-13  # absolute floor for log-densities that happen to be ~0 (|log p| < 1e-3)


def assert_parity(got, want, rtol, atol=0.0, what=""):
    got = np.atleast_1d(np.asarray(got, dtype=np.float64))
    want = np.atleast_1d(np.asarray(want, dtype=np.float64))
    assert got.shape == want.shape, (what, got.shape, want.shape)
    nan_g, nan_w = np.isnan(got), np.isnan(want)
    assert np.array_equal(nan_g, nan_w), "%s: NaN classes differ at %s" % (what, np.where(nan_g != nan_w)[0][:10])
    inf_w = np.isinf(want)
    assert np.array_equal(got[inf_w], want[inf_w]), "%s: inf classes differ" % what
    assert not np.isinf(got[~inf_w & ~nan_w]).any(), "%s: unexpected inf" % what
    zero_w = want == 0
    assert np.array_equal(got == 0, zero_w), "%s: exact-zero classes differ at %s" % (
        what, np.where((got == 0) != zero_w)[0][:10])
    m = ~(nan_w | inf_w | zero_w)
    if m.any():
        err = np.abs(got[m] - want[m])
        # below DBL_MIN the reference's own result carries an absolute rounding error of one
        # subnormal spacing per operation (src/pdf.pxi:102 rounds three times there)
        lim = rtol * np.abs(want[m]) + atol + 8 * 4.9406564584124654e-324
        worst = np.argmax(err - lim)
        assert (err <= lim).all(), "%s: max rel err %.3e (limit %.1e) at %d: got %r want %r" % (
            what, np.max(err / np.abs(want[m])), rtol, np.where(m)[0][worst], got[m][worst], want[m][worst])
    return float(np.max(np.abs(got[m] - want[m]) / np.abs(want[m]))) if m.any() else 0.0


PNAMES = ("v", "sv", "a", "z", "sz", "t", "st")


def draw_params(rng, n, sv_on=True, sz_on=True, st_on=True):
    """Ranges of hddm/generate.py:38-46; draw order of SURVEY.md 8(d): sv, sz, st, z, v, t, a."""
    sv = rng.uniform(0, 2.5, n) if sv_on else np.zeros(n)
    sz = rng.uniform(0, 0.4, n) if sz_on else np.zeros(n)
    st = rng.uniform(0, 0.35, n) if st_on else np.zeros(n)
    z = rng.uniform(0.4, 0.6, n)
    v = rng.uniform(-4, 4, n)
    t = rng.uniform(0.2, 0.5, n)
    a = rng.uniform(0.5, 2.0, n)
    return np.stack([v, sv, a, z, sz, t, st], axis=0)  # [7, n]


def draw_rts(rng, n, onset):
    g = rng.gamma(2.0, 0.25, n)
    sgn = np.where(rng.random(n) < 0.6, 1.0, -1.0)
    return (onset + g) * sgn


def sweep_inputs(T, P, onset=0.70, seed=20261104, p_outlier=0.05, w_outlier=0.1):
    """The C4 sweep of SURVEY.md 8(d): S-A onset 0.70, S-B onset 0.30.  Returns x[T], params[9,P]."""
    rng = np.random.default_rng(seed)
    par7 = draw_params(rng, P)
    x = draw_rts(rng, T, onset)
    params = np.concatenate([par7, np.full((1, P), p_outlier), np.full((1, P), w_outlier)], axis=0)
    return x, np.ascontiguousarray(params)


def cdf_noise_band(O, x, p, po, wo):
    """Conditioning of the reference's own dmat_cdf_array at these inputs: the largest change of its
    output under a ONE-ulp change of v, a, z or t.  With sz or st below ~1e-3 the wrapper substitutes
    1e-10 epsilons (src/cdfdif_wrapper.pyx:36-42) and cdfdif differences series evaluated 1e-10 apart
    (src/cdfdif.c:135-140,148), so the reference result itself is only defined to ~1e-7 relative there;
    for the full DDM (sz, st >= 1e-2) the band is ~1e-14 and the 1e-10 bar is the binding one."""
    y0 = O.dmat_cdf_array(x, *p, po, wo)
    band = np.zeros_like(y0)
    for j in (0, 2, 3, 5):
        for s in (1.0, -1.0):
            q = np.array(p, dtype=np.float64)
            q[j] = np.nextafter(q[j], q[j] + s)
            try:
                band = np.maximum(band, np.abs(O.dmat_cdf_array(x, *q, po, wo) - y0))
            except ValueError:
                pass
    return band
