"""CPU tests that PIN the oracle (oracle/wfpt_oracle.c, oracle/cdfdif_oracle.c):
  1. against the reference's own MATLAB golden vectors (hddm/tests/matlab_values.py:3-162);
  2. bit for bit against the committed outputs of the compiled unmodified reference
     (tests/golden/ref_*.npz, made by tests/golden/make_golden.py from oracle/_ref);
  3. bit for bit against oracle/_ref live, when that build is present (this container);
  4. through the reference's own analytic / cross-check tests (hddm/tests/test_likelihoods.py),
     re-hosted on pytest with the kabuki-dependent parameter draws inlined.
"""
import json
import os

import numpy as np
import pytest
import scipy.integrate

from oracle import oracle as O
from oracle import ref
from tests.util import draw_params, draw_rts

simps = getattr(scipy.integrate, "simps", None) or scipy.integrate.simpson


def same(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    return np.array_equal(a, b, equal_nan=True)


def ulps(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    m = np.isfinite(a) & np.isfinite(b)
    assert np.array_equal(a[~m], b[~m], equal_nan=True)
    return np.max(np.abs(a[m] - b[m]) / np.spacing(np.abs(b[m]))) if m.any() else 0.0


def test_matlab_golden_vectors(golden_dir):
    vals = json.load(open(os.path.join(golden_dir, "matlab_wfpt.json")))["vals"]
    assert len(vals) == 20
    worst = 0.0
    for r in vals:
        got = O.full_pdf(-r["rt"], r["v"], 0, r["a"], r["z"], 0, r["t"], 0, r["err"], 0)
        np.testing.assert_array_almost_equal(r["matlab_wfpt"], got, 9)   # the reference's own criterion
        worst = max(worst, abs(got - r["matlab_wfpt"]))
    assert worst < 1e-15   # survey probe: the compiled reference sits at 5.6e-17


def test_full_pdf_fixture_bitwise(golden_dir):
    d = np.load(os.path.join(golden_dir, "ref_full_pdf.npz"))
    rows, want = d["rows"], d["out"]
    got = np.array([O.full_pdf(r[0], *r[1:8], r[8], int(r[9]), int(r[10]), int(r[11]), r[12]) for r in rows])
    assert same(got, want)
    assert (want == 0).sum() >= 24     # failure-mode zeros, hddm/tests/test_likelihoods.py:225-262


def test_likelihood_fixtures_bitwise(golden_dir):
    d = np.load(os.path.join(golden_dir, "ref_likelihoods.npz"))
    for k in range(int(d["wl_n"])):
        p = d["wl_p_%d" % k]
        got = O.wiener_like(d["wl_x_%d" % k], *p[:7], p[7], int(p[8]), int(p[9]), int(p[10]), p[11], p[12], p[13])
        assert same(got, d["wl_y_%d" % k]), k
    assert O.wiener_like(np.array([1., 2., 3., 0.]), 1, 0, 2, .5, 0, 0, 0, 1e-4) == -np.inf
    for k in range(int(d["pa_n"])):
        p = d["pa_p_%d" % k]
        with np.errstate(divide="ignore"):
            got = O.pdf_array(d["pa_x_%d" % k], *p[:7], p[7], int(p[8]), int(p[9]), int(p[10]), int(p[11]), p[12],
                              p[13], p[14])
        if int(p[8]):   # the reference logs with numpy's vectorised np.log: allow its last-ulp difference
            assert ulps(got, d["pa_y_%d" % k]) <= 4
        else:
            assert same(got, d["pa_y_%d" % k]), k
    names = ("v", "sv", "a", "z", "sz", "t", "st")
    for k in range(int(d["wm_n"])):
        cols, multi = d["wm_cols_%d" % k], [names[j] for j in d["wm_multi_%d" % k]]
        args = [cols[:, j].copy() if names[j] in multi else cols[0, j] for j in range(7)]
        got = O.wiener_like_multi(d["wm_x_%d" % k], *args, 1e-4, set(multi), 10, 10, 1, 1e-3, 0.05, 0.1)
        assert same(got, d["wm_y_%d" % k]), k
    for k in range(int(d["wc_n"])):
        got = O.wiener_like_contaminant(d["wc_x_%d" % k], d["wc_c_%d" % k], *d["wc_p_%d" % k], 0.1, 5.0, 1e-4,
                                        10, 10, 1, 1e-6)
        assert same(got, d["wc_y_%d" % k]), k


def test_cdf_fixtures_bitwise(golden_dir):
    d = np.load(os.path.join(golden_dir, "ref_cdf.npz"))
    for k in range(int(d["cd_n"])):
        p = d["cd_p_%d" % k]
        assert same(O.dmat_cdf_array(d["cd_x_%d" % k], *p[:7], p[7], p[8]), d["cd_y_%d" % k]), k
    for j in range(int(d["gc_n"])):
        x, c = O.gen_cdf_using_pdf(*d["gc_p_%d" % j], 1e-4, 500, 5., 2, 2, 1, 1e-3, 0.05, 0.1)
        assert same(x, d["gc_x_%d" % j]) and same(c, d["gc_y_%d" % j])
    for j in range(int(d["gr_n"])):
        got = O.gen_rts_from_cdf(*d["gr_p_%d" % j], 400, -4.0, 4.0, 1e-3, f=d["gr_f_%d" % j], u_delay=d["gr_u_%d" % j])
        assert same(got, d["gr_y_%d" % j])


def test_cdf_guards():
    x = np.array([0.5, -0.7, 6.0])
    with pytest.raises(AssertionError):
        O.dmat_cdf_array(x, 1., .5, 2., .5, .2, .3, .1, 0.05, 0.1)
    with pytest.raises(ValueError):
        O.dmat_cdf_array(x, 1., .5, 2., .95, .2, .3, .1, 0.0, 0.1)
    with pytest.raises(ValueError):
        O.gen_cdf_using_pdf(1., .5, 2., .95, .2, .3, .1, 1e-4)


@pytest.mark.skipif(not ref.available(), reason="oracle/_ref not built (needs /root/reference)")
def test_live_against_compiled_reference():
    W, CW = ref.wfpt(), ref.cdfdif_wrapper()
    rng = np.random.default_rng(31)
    n_bad = 0
    for (nn, se, ad) in [(2, 1e-3, 1), (10, 1e-3, 1), (6, 1e-6, 1), (4, 0.0, 0)]:
        for q in range(8):
            P = draw_params(rng, 40, q & 1, q & 2, q & 4)
            x = draw_rts(rng, 40, 0.30)
            err = 10 ** rng.uniform(-10, -2, 40)
            for i in range(40):
                r = W.full_pdf(x[i], *P[:, i], err[i], nn, nn, ad, se)
                o = O.full_pdf(x[i], *P[:, i], err[i], nn, nn, ad, se)
                n_bad += not same(r, o)
    assert n_bad == 0
    x = draw_rts(rng, 500, 0.4)
    p = draw_params(rng, 1)[:, 0]
    assert W.wiener_like(x, *p, 1e-4, 2, 2, 1, 1e-3, .05, .1) == O.wiener_like(x, *p, 1e-4, 2, 2, 1, 1e-3, .05, .1)
    assert same(CW.dmat_cdf_array(x, *p, .05, .1), O.dmat_cdf_array(x, *p, .05, .1))


# ---- the reference's own tests, re-hosted (hddm/tests/test_likelihoods.py) -----------------------

def test_summed_logp():
    """:79-97 -- sum(pdf_array(logp)) ~ wiener_like; and the rt=0 -> -inf case."""
    rng = np.random.default_rng(3123)
    for q in (0, 1, 3, 7):
        p = draw_params(rng, 1, q & 1, q & 2, q & 4)[:, 0]
        x = draw_rts(rng, 300, 0.7)
        summed = np.sum(O.pdf_array(x, *p, 1e-4, logp=1, n_st=10, n_sz=10, simps_err=1e-8))
        np.testing.assert_almost_equal(summed, O.wiener_like(x, *p, 1e-4), 8)
    assert O.wiener_like(np.array([1., 2., 3., 0.]), 1, 0, 2, .5, 0, 0, 0, 1e-4) == -np.inf


def test_pdf_sv_against_quadrature():
    """:99-114 -- analytic sv marginal vs numerical integration of the sv=0 density over v."""
    from scipy.stats import norm
    rng = np.random.default_rng(5)
    for _ in range(12):
        sv, v, t, a, z = rng.uniform(.1, 2.), (rng.random() - .5) * 4, rng.random() * .5, 1.5 + rng.random(), .5 * rng.random() + .25
        rt = rng.random() * 4 + t
        err = 10 ** (-3 - np.ceil(rng.random() * 12))
        f = lambda vv: O.full_pdf(rt, vv, 0, a, z, 0, t, 0, err) * norm.pdf(vv, v, sv)
        numeric = scipy.integrate.quad(f, v - 12 * sv, v + 12 * sv, epsabs=1e-12, epsrel=1e-12, limit=400)[0]
        np.testing.assert_almost_equal(O.full_pdf(rt, v, sv, a, z, 0, t, 0, err), numeric, 6)


def test_adaptive_vs_fixed():
    """:117-144 -- adaptive (n=5) vs 60-interval fixed Simpson over st, 3 decimals."""
    rng = np.random.default_rng(6)
    for _ in range(20):
        sv, v, t, a, z = rng.uniform(0, 2.), (rng.random() - .5) * 4, .3 + rng.random() * .2, 1.5 + rng.random(), .4 + .2 * rng.random()
        st = rng.uniform(.05, .3)
        rt = (t + st / 2 + rng.random() * 2) * (1 if rng.random() < .5 else -1)
        fixed = O.full_pdf(rt, v, sv, a, z, 0, t, st, 1e-10, 60, 0, 0)
        adapt = O.full_pdf(rt, v, sv, a, z, 0, t, st, 1e-10, 5, 5, 1, 1e-8)
        np.testing.assert_almost_equal(fixed, adapt, 3)


def test_pdf_integrates_to_one():
    """:147-161."""
    rng = np.random.default_rng(123)
    for q in (0, 1, 7):
        p = draw_params(rng, 1, q & 1, q & 2, q & 4)[:, 0]
        f = lambda xx: np.exp(O.wiener_like(np.array([xx]), *p, 1e-4, 2, 2, 1, 1e-3))
        total = scipy.integrate.quad(f, -8, 0, limit=300)[0] + scipy.integrate.quad(f, 0, 8, limit=300)[0]
        np.testing.assert_almost_equal(total, 1.0, 2)


def test_fixed_simpson_against_grid():
    """:163-222 -- full_pdf with sz / st / both vs scipy Simpson over point evaluations (2 decimals)."""
    rng = np.random.default_rng(8)
    for _ in range(6):
        sv, v, t, a, z = rng.uniform(0, 1.), (rng.random() - .5) * 3, .35 + rng.random() * .1, 1.5 + rng.random(), .5
        sz, st = rng.uniform(.05, .3), rng.uniform(.05, .3)
        rt = -(t + st / 2 + .2 + rng.random())
        n = 60
        zs = np.linspace(z - sz / 2, z + sz / 2, n + 1)
        ts = np.linspace(t - st / 2, t + st / 2, n + 1)
        over_z = simps([O.full_pdf(rt, v, sv, a, zz, 0, t, 0, 1e-10) for zz in zs], x=zs) / sz
        np.testing.assert_almost_equal(O.full_pdf(rt, v, sv, a, z, sz, t, 0, 1e-10, 0, n, 0), over_z, 6)
        over_t = simps([O.full_pdf(rt, v, sv, a, z, 0, tt, 0, 1e-10) for tt in ts], x=ts) / st
        np.testing.assert_almost_equal(O.full_pdf(rt, v, sv, a, z, 0, t, st, 1e-10, n, 0, 0), over_t, 6)
        both = simps([simps([O.full_pdf(rt, v, sv, a, zz, 0, tt, 0, 1e-10) for zz in zs], x=zs) / sz for tt in ts], x=ts) / st
        np.testing.assert_almost_equal(O.full_pdf(rt, v, sv, a, z, sz, t, st, 1e-10, n, n, 0), both, 6)
        np.testing.assert_almost_equal(O.full_pdf(rt, v, sv, a, z, sz, t, st, 1e-10, 10, 10, 1, 1e-8), both, 2)


def test_failure_modes_are_exact_zero():
    """:225-262."""
    ok = dict(v=1., sv=.3, a=2., z=.5, sz=.2, t=.3, st=.1)
    order = ("v", "sv", "a", "z", "sz", "t", "st")
    for bad in (dict(z=-.1), dict(z=1.1), dict(z=.05), dict(z=.95), dict(a=-1.), dict(t=-.1), dict(st=-.1),
                dict(sv=-.1), dict(sz=-.1), dict(sz=1.1), dict(t=.02), dict(t=1.5)):
        p = dict(ok, **bad)
        for x in (.9, -.9):
            assert O.full_pdf(x, *[p[k] for k in order], 1e-4) == 0.0


def test_eval_counter_matches_survey():
    """the adaptive 2-D recursion makes 24.9 pdf_sv calls per eval on sweep S-A."""
    from tests.util import sweep_inputs
    x, params = sweep_inputs(600, 10)
    O.pdfsv_calls(reset=True)
    O.wiener_like_batch(x, params, err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3)
    mean = O.pdfsv_calls(reset=True) / 6000.0
    assert 20.0 < mean < 27.0, mean   # 10 sets only: a near-zero sz or st draw lowers it


# ---- the Python-side oracles, pinned on outputs of the reference's own functions ----

def test_ppc_oracle_reproduces_reference_gen_ppc_stats(golden_dir):
    """oracle/ppc_oracle.py against tests/golden/ref_ppc_stats.npz: the fifteen lambdas of gen_ppc_stats()
    (hddm/utils.py:241-267), executed from the unmodified reference file by tests/golden/make_golden_py.py, on
    19 data sets (empty boundary, ties, zeros, single RTs, zero variance).  Bit for bit, NaN classes included."""
    from oracle.ppc_oracle import ppc_stats
    d = np.load(os.path.join(golden_dir, "ref_ppc_stats.npz"))
    assert [str(n) for n in d["names"]] == ["accuracy", "mean_ub", "std_ub", "10q_ub", "30q_ub", "50q_ub", "70q_ub",
                                            "90q_ub", "mean_lb", "std_lb", "10q_lb", "30q_lb", "50q_lb", "70q_lb",
                                            "90q_lb"]
    assert int(d["n"]) >= 19
    nan_rows = 0
    for k in range(int(d["n"])):
        want = d["y_%d" % k]
        assert same(ppc_stats(d["x_%d" % k]), want), k
        nan_rows += bool(np.isnan(want).any())
    assert nan_rows >= 6


def simdrift_cases(golden_dir):
    d = np.load(os.path.join(golden_dir, "ref_simdrift.npz"))
    keys = [str(k) for k in d["keys"]]
    for k in range(int(d["n"])):
        cfg = d["cfg_%d" % k]
        params = {n: float(v) for n, v in zip(keys, d["par_%d" % k]) if not np.isnan(v)}  # NaN: key absent
        yield k, params, int(cfg[0]), float(cfg[1]), float(cfg[2]), d["u_init_%d" % k], d["u_steps_%d" % k], \
            d["rts_%d" % k], d["blocks_%d" % k]


def test_simulator_oracle_reproduces_reference_runs(golden_dir):
    """oracle/sim_oracle.py against tests/golden/ref_simdrift.npz: ten runs of _gen_rts_from_simulated_drift
    (hddm/generate.py:208-319, executed unmodified by make_golden_py.py with numpy's global RNG seeded and the
    consumed random numbers recorded), four of them with v_switch.  Bit for bit."""
    from oracle import sim_oracle
    n, switched, multi_block = 0, 0, 0
    for k, params, S, dt, intra_sv, u_init, u_steps, want, blocks in simdrift_cases(golden_dir):
        got = sim_oracle.gen_rts_from_simulated_drift(params, u_init, u_steps, dt=dt, intra_sv=intra_sv)
        assert same(got, want), k
        assert not np.isnan(want).any()
        n += 1
        switched += "v_switch" in params
        multi_block += bool((blocks > 1).any())
    assert n >= 10 and switched >= 4 and multi_block >= 6
