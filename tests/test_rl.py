"""Reinforcement-learning likelihoods (src/wfpt.pyx:79-241, 277-320).

CPU part: the C oracle reproduces, bit for bit, the outputs of the compiled unmodified reference stored
in tests/golden/ref_rl.npz (the reference has no known-answer tests for these functions, so this is
their pin).  GPU part: the CUDA path through the reference-shaped API -> C ABI against the fixtures
and against the oracle on fresh seeded tasks, at the north_star's 1e-9 relative summed log-likelihood."""
import os

import numpy as np
import pytest

from tests.util import RTOL_SUM, assert_parity, draw_params

RLN = ("v", "sv", "a", "z", "sz", "t", "st", "alpha")


@pytest.fixture(scope="module")
def O():
    from oracle import oracle
    oracle.build()
    return oracle


@pytest.fixture(scope="module")
def fx(golden_dir):
    return np.load(os.path.join(golden_dir, "ref_rl.npz"))


def _case(fx, k):
    p = fx["rl_p_%d" % k]
    return dict(x=fx["rl_x_%d" % k], response=fx["rl_resp_%d" % k], feedback=fx["rl_fb_%d" % k],
                split_by=fx["rl_split_%d" % k], q=p[0], alpha=p[1], pos_alpha=p[2], ddm=p[3:10],
                num=(p[10], int(p[11]), int(p[12]), int(p[13]), p[14]), outl=(p[15], p[16]))


def _multi_args(fx, k):
    cols, multi = fx["rl_mcols_%d" % k], [RLN[j] for j in fx["rl_multi_%d" % k]]
    return [np.ascontiguousarray(cols[:, j]) if RLN[j] in multi else cols[0, j] for j in range(8)], multi


def _all_three(M, fx, k):
    c = _case(fx, k)
    ddm = M.wiener_like_rlddm(c["x"], c["response"], c["feedback"], c["split_by"], c["q"], c["alpha"],
                              c["pos_alpha"], *c["ddm"], *c["num"], *c["outl"])
    choice = M.wiener_like_rl(c["response"], c["feedback"], c["split_by"], c["q"], c["alpha"], c["pos_alpha"],
                              c["ddm"][0], c["ddm"][3], p_outlier=c["outl"][0], w_outlier=c["outl"][1])
    args, multi = _multi_args(fx, k)
    reg = M.wiener_like_multi_rlddm(c["x"], c["response"], c["feedback"], fx["rl_msplit_%d" % k], c["q"], *args,
                                    1e-4, multi, 2, 2, 1, 1e-3, *c["outl"])
    return ddm, choice, reg


def test_oracle_reproduces_reference_fixture_bitwise(O, fx):
    for k in range(int(fx["rl_n"])):
        ddm, choice, reg = _all_three(O, fx, k)
        assert ddm == float(fx["rl_ddm_%d" % k]), k
        assert choice == float(fx["rl_choice_%d" % k]), k
        assert reg == float(fx["rl_my_%d" % k]), k


def test_oracle_live_against_compiled_reference(O):
    from oracle import ref
    if not ref.available():
        pytest.skip("oracle/_ref not built here (needs /root/reference)")
    W = ref.wfpt()
    rng = np.random.default_rng(31)
    for k in range(8):
        x, response, feedback, split_by = _task(rng, 150, 1 + k % 4)
        p = draw_params(rng, 1)[:, 0]
        p[0] = abs(p[0]) + 0.5
        a = (x, response, feedback, split_by, 0.5, rng.normal(), 100.0 if k % 2 else rng.normal(), *p, 1e-4)
        kw = dict(n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, p_outlier=0.05, w_outlier=0.1)
        assert O.wiener_like_rlddm(*a, **kw) == W.wiener_like_rlddm(*a, **kw)


def _task(rng, n, n_cond):
    split_by = rng.integers(0, n_cond, n).astype(np.int64) * 3 - 2  # labels need not be 0..C-1
    response = rng.integers(0, 2, n).astype(np.int64)
    feedback = (rng.random(n) < np.where(response == 1, 0.8, 0.3)).astype(np.float64)
    x = (0.35 + rng.gamma(2.0, 0.25, n)) * np.where(response == 1, 1.0, -1.0)
    return x, response, feedback, split_by


def test_oracle_rl_edge_cases(O):
    e, l = np.zeros(0), np.zeros(0, dtype=np.int64)
    assert O.wiener_like_rlddm(e, l, e, l, .5, 0., 100., 1., 0, 2., .5, 0, .3, 0, 1e-4) == 0.0
    assert O.wiener_like_rl(l, e, l, .5, 0., 100., 1., .5) == 0.0
    x, r, f, s = _task(np.random.default_rng(2), 40, 2)
    assert O.wiener_like_rlddm(x, r, f, s, .5, 0., 100., 1., 0, 2., .5, 0, .3, 0, 1e-4, p_outlier=1.5) == -np.inf
    # one condition of a single trial contributes nothing: only the Q update runs (src/wfpt.pyx:120-134)
    one = O.wiener_like_rlddm(x[:1], r[:1], f[:1], s[:1], .5, 0., 100., 1., 0, 2., .5, 0, .3, 0, 1e-4)
    assert one == 0.0


# ---- GPU ---------------------------------------------------------------------------------------

@pytest.fixture(scope="module")
def W(gpu):
    from hddm_b200 import wfpt
    return wfpt


@pytest.mark.gpu
def test_gpu_against_reference_fixture(W, fx):
    worst = 0.0
    for k in range(int(fx["rl_n"])):
        got = _all_three(W, fx, k)
        want = (float(fx["rl_ddm_%d" % k]), float(fx["rl_choice_%d" % k]), float(fx["rl_my_%d" % k]))
        worst = max(worst, assert_parity(got, want, RTOL_SUM, what="rl fixture %d" % k))
    print("rl fixtures: max rel err %.2e" % worst)


@pytest.mark.gpu
@pytest.mark.parametrize("kernel", [0, 1, 2])
def test_gpu_rlddm_against_oracle(W, O, kernel):
    from hddm_b200 import _lib
    _lib.check(_lib.lib().hddm_b200_set_kernel(kernel))
    try:
        rng = np.random.default_rng(77 + kernel)
        for k in range(6):
            n = (5000, 333, 2049, 64, 1, 700)[k]
            x, response, feedback, split_by = _task(rng, n, (3, 1, 4, 2, 1, 5)[k])
            p = draw_params(rng, 1, k & 1, k & 2, k & 4)[:, 0]
            p[0] = abs(p[0]) + 0.5
            q, alpha, pos_alpha = 0.5, rng.normal(0, 1.5), (100.0 if k % 2 else rng.normal(0, 1.5))
            kw = dict(n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, p_outlier=0.05, w_outlier=0.1)
            got = W.wiener_like_rlddm(x, response, feedback, split_by, q, alpha, pos_alpha, *p, 1e-4, **kw)
            want = O.wiener_like_rlddm(x, response, feedback, split_by, q, alpha, pos_alpha, *p, 1e-4, **kw)
            assert_parity(got, want, RTOL_SUM, what="rlddm case %d" % k)
            got = W.wiener_like_rl(response, feedback, split_by, q, alpha, pos_alpha, p[0], p[3], p_outlier=0.05,
                                   w_outlier=0.1)
            want = O.wiener_like_rl(response, feedback, split_by, q, alpha, pos_alpha, p[0], p[3], p_outlier=0.05,
                                    w_outlier=0.1)
            assert_parity(got, want, RTOL_SUM, what="rl case %d" % k)
    finally:
        _lib.check(_lib.lib().hddm_b200_set_kernel(0))


@pytest.mark.gpu
def test_gpu_rl_batch_equals_scalar_calls_and_oracle(W, O):
    """The batched entry points (many proposals against one subject's task in one launch pair) give,
    per set, the value of the scalar call -- and the pdf_sv evaluation count of the oracle, i.e. the same
    series / Simpson decisions."""
    rng = np.random.default_rng(5)
    x, response, feedback, split_by = _task(rng, 900, 3)
    P = 37
    par7 = draw_params(rng, P)
    par7[0] = np.abs(par7[0]) + 0.5
    params = np.concatenate([par7, np.full((1, P), 0.05), np.full((1, P), 0.1), np.full((1, P), 0.5),
                             rng.normal(0, 1.5, (1, P)), np.where(rng.random((1, P)) < 0.5, 100.0, rng.normal(0, 1.5, (1, P)))])
    nm = dict(err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3)
    got, evals = W.wiener_like_rlddm_batch(x, response, feedback, split_by, params, return_evals=True, **nm)
    O.pdfsv_calls(reset=True)
    want = np.array([O.wiener_like_rlddm(x, response, feedback, split_by, params[9, p], params[10, p], params[11, p],
                                         *params[:7, p], nm["err"], nm["n_st"], nm["n_sz"], 1, nm["simps_err"],
                                         params[7, p], params[8, p]) for p in range(P)])
    want_evals = O.pdfsv_calls(reset=True)
    assert_parity(got, want, RTOL_SUM, what="rlddm batch")
    assert evals == want_evals
    one = np.array([W.wiener_like_rlddm(x, response, feedback, split_by, params[9, p], params[10, p], params[11, p],
                                        *params[:7, p], nm["err"], nm["n_st"], nm["n_sz"], 1, nm["simps_err"],
                                        params[7, p], params[8, p]) for p in range(P)])
    assert np.array_equal(one, got)  # deterministic: batch of P == P scalar launches, bit for bit
    cp = np.stack([params[9], params[10], params[11], params[0], params[3], params[7], params[8]])
    gotc = W.wiener_like_rl_batch(response, feedback, split_by, cp)
    wantc = np.array([O.wiener_like_rl(response, feedback, split_by, *cp[:5, p], p_outlier=cp[5, p], w_outlier=cp[6, p])
                      for p in range(P)])
    assert_parity(gotc, wantc, RTOL_SUM, what="rl batch")


@pytest.mark.gpu
def test_gpu_multi_rlddm_against_oracle(W, O):
    rng = np.random.default_rng(9)
    for k, multi in enumerate([("v",), ("alpha",), ("v", "alpha", "a"), ("z", "sz"), RLN]):
        n = 400 + 300 * k
        x, response, feedback, split_by = _task(rng, n, 3)
        split_by = np.sort(split_by) if k % 2 else split_by  # unsorted labels: a reset at every change
        base = np.append(draw_params(rng, 1)[:, 0], rng.normal())
        base[0] = abs(base[0]) + 0.5
        per = np.concatenate([draw_params(rng, n), rng.normal(0, 1, (1, n))])
        per[0] = np.abs(per[0]) + 0.5
        args = [np.ascontiguousarray(per[j]) if RLN[j] in multi else base[j] for j in range(8)]
        kw = dict(n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3, p_outlier=0.05, w_outlier=0.1)
        got = W.wiener_like_multi_rlddm(x, response, feedback, split_by, 0.5, *args, 1e-4, list(multi), **kw)
        want = O.wiener_like_multi_rlddm(x, response, feedback, split_by, 0.5, *args, 1e-4, list(multi), **kw)
        assert_parity(got, want, RTOL_SUM, what="multi_rlddm %s" % (multi,))


@pytest.mark.gpu
def test_gpu_rl_edge_cases(W):
    e, l = np.zeros(0), np.zeros(0, dtype=np.int64)
    assert W.wiener_like_rlddm(e, l, e, l, .5, 0., 100., 1., 0, 2., .5, 0, .3, 0, 1e-4) == 0.0
    assert W.wiener_like_rl(l, e, l, .5, 0., 100., 1., .5) == 0.0
    x, r, f, s = _task(np.random.default_rng(2), 40, 2)
    assert W.wiener_like_rlddm(x, r, f, s, .5, 0., 100., 1., 0, 2., .5, 0, .3, 0, 1e-4, p_outlier=1.5) == -np.inf
    assert W.wiener_like_rl(r, f, s, .5, 0., 100., 1., .5, p_outlier=-0.5) == -np.inf
    assert W.wiener_like_rlddm(x[:1], r[:1], f[:1], s[:1], .5, 0., 100., 1., 0, 2., .5, 0, .3, 0, 1e-4) == 0.0
    xz = x.copy()
    xz[7] = np.sign(xz[7]) * 0.01  # before the onset: density 0 -> -inf without the outlier floor
    assert W.wiener_like_rlddm(xz, r, f, s, .5, 0., 100., 1., 0, 2., .5, 0, .3, 0, 1e-4) == -np.inf
    with pytest.raises(ValueError):
        W.wiener_like_rlddm(x, r.astype(np.int32), f, s, .5, 0., 100., 1., 0, 2., .5, 0, .3, 0, 1e-4)
    with pytest.raises(ValueError):
        W.wiener_like_rlddm(x, r + 1, f, s, .5, 0., 100., 1., 0, 2., .5, 0, .3, 0, 1e-4)


@pytest.mark.gpu
def test_gpu_rl_stochastics_on_dataframes(W, O):
    """The logp closures of hddm/models/hddm_rl.py:61-73, hddm/models/rl.py:72-85 and
    hddm/models/hddm_rl_regression.py:25-38 on a DataFrame-valued observed node."""
    import pandas as pd
    from hddm_b200 import likelihoods as L
    x, response, feedback, split_by = _task(np.random.default_rng(4), 300, 3)
    df = pd.DataFrame(dict(rt=x, response=response.astype(float), feedback=feedback, split_by=split_by, q_init=0.5))
    par = dict(v=2.0, alpha=-0.5, pos_alpha=0.3, sv=0.4, a=1.6, z=0.5, sz=0.1, t=0.2, st=0.1)
    node = L.WienerRL(value=df, p_outlier=0.05, **par)
    want = O.wiener_like_rlddm(x, response, feedback, split_by, 0.5, par["alpha"], par["pos_alpha"], par["v"], par["sv"],
                               par["a"], par["z"], par["sz"], par["t"], par["st"], 1e-4, 2, 2, 1, 1e-3, 0.05, 0.1)
    assert_parity(node.logp, want, RTOL_SUM, what="WienerRL.logp")
    node = L.RL(value=df, v=2.0, alpha=-0.5, pos_alpha=100.0, z=0.5, p_outlier=0.05)
    want = O.wiener_like_rl(response, feedback, split_by, 0.5, -0.5, 100.0, 2.0, 0.5, p_outlier=0.05, w_outlier=0.1)
    assert_parity(node.logp, want, RTOL_SUM, what="RL.logp")
    vcol = pd.Series(np.random.default_rng(5).uniform(1, 3, 300), index=df.index)
    got = L.wienerRL_multi_like(df, vcol, 0.4, 1.6, 0.5, 0.1, 0.2, 0.1, -0.5, ['v'], p_outlier=0.05)
    want = O.wiener_like_multi_rlddm(x, response, feedback, split_by, 0.5, vcol.values, 0.4, 1.6, 0.5, 0.1, 0.2, 0.1,
                                     -0.5, 1e-4, ['v'], p_outlier=0.05, w_outlier=0)
    assert_parity(got, want, RTOL_SUM, what="wienerRL_multi_like")
