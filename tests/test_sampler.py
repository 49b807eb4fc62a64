"""Host logic of the batched sampler (hddm_b200/sampler.py) and of the likelihood glue
(hddm_b200/likelihoods.py) that needs no GPU: slice steps against known targets, the model's update
order with a stand-in likelihood, pickling without device handles."""
import pickle

import numpy as np

from hddm_b200 import sampler


def test_slice_step_samples_a_normal():
    rng = np.random.default_rng(1)
    mu, sd = np.array([0.0, 3.0, -2.0]), np.array([1.0, 0.5, 2.0])
    logp = lambda x: -0.5 * ((x - mu) / sd) ** 2
    x = np.zeros(3)
    draws = []
    for _ in range(4000):
        x, lp, calls = sampler.slice_step(x, logp, 1.0, rng)
        assert calls >= 4
        draws.append(x)
    d = np.array(draws[200:])
    assert np.allclose(d.mean(0), mu, atol=0.12)
    assert np.allclose(d.std(0), sd, rtol=0.12)


def test_slice_step_respects_left_bound_and_shape():
    rng = np.random.default_rng(2)
    logp = lambda x: np.where(x > 0, -x, -np.inf)  # Exp(1) on each element
    x = np.full((2, 3), 0.5)
    draws = []
    for _ in range(3000):
        x, _, _ = sampler.slice_step(x, logp, 1.0, rng, left=0.0)
        assert x.shape == (2, 3) and (x > 0).all()
        draws.append(x)
    assert abs(np.mean(draws) - 1.0) < 0.08


def test_prior_densities_match_scipy():
    from scipy import stats
    x = np.array([0.2, 1.0, 2.5])
    assert np.allclose(sampler._gamma_from_mean_std(x, 1.5, 0.75), stats.gamma.logpdf(x, 4.0, scale=1 / (1.5 / 0.75 ** 2)))
    assert np.allclose(sampler._halfnorm_logpdf(x, 2.0), stats.halfnorm.logpdf(x, scale=2.0))
    assert np.allclose(sampler._beta_logpdf(x[:1], 1.0, 3.0), stats.beta.logpdf(x[:1], 1, 3))
    assert np.allclose(sampler._norm_logpdf(x, 2.0, 3.0), stats.norm.logpdf(x, 2.0, 3.0))
    assert sampler._gamma_from_mean_std(np.array([-1.0]), 1.0, 1.0)[0] == -np.inf


def test_hierarchical_sweep_with_standin_likelihood():
    """The update order touches the likelihood only for subject-level and group-only parameters, always
    with chains x subjects sets, and chains are independent and reproducible."""
    S, C = 5, 3
    seen = []

    def loglik(set_node, params):
        assert params.shape == (9, C * S) and np.array_equal(set_node, np.tile(np.arange(S), C))
        seen.append(params.copy())
        v, sv, a, sz, t, st = params[0], params[1], params[2], params[4], params[5], params[6]
        return -((v - 1.0) ** 2 + (a - 1.2) ** 2 * 4 + (t - 0.3) ** 2 * 50 + (sv - 1) ** 2 + (sz - .2) ** 2 + (st - .1) ** 2)

    m = sampler.HierarchicalDDM(S, loglik, chains=C, seed=7, batched=False)  # one launch of C x S sets per probe
    tr, dt = m.sample(5)
    assert tr['v_subj'].shape == (5, C, S) and tr['sv'].shape == (5, C)
    assert m.loglik_calls == len(seen) and m.sets_evaluated == len(seen) * C * S
    assert np.isfinite(tr['a_subj']).all() and (tr['a_subj'] > 0).all() and (tr['t_subj'] > 0).all()
    assert (tr['sv'] > 0).all() and ((tr['sz'] > 0) & (tr['sz'] < 1)).all()
    m2 = sampler.HierarchicalDDM(S, loglik, chains=C, seed=7, batched=False)
    tr2, _ = m2.sample(5)
    assert all(np.array_equal(tr[k], tr2[k]) for k in tr)


def test_batched_slice_step_equals_the_sequential_one():
    """slice_step_batched evaluates speculative probes in a few large batches: same draws, same decisions, same
    generator state afterwards as slice_step (with and without a left bound)."""
    mu, sd = np.array([0.0, 3.0, -2.0, 1.0]), np.array([1.0, 0.5, 2.0, 0.05])
    for left in (None, -1.0):
        base = lambda x: -0.5 * ((x - mu) / sd) ** 2
        f = (lambda x: np.where(x > left, base(x), -np.inf)) if left is not None else base
        r1, r2 = np.random.default_rng(5), np.random.default_rng(5)
        x = np.array([0.1, 2.0, 0.5, 0.9])
        xb, lpb, calls, batches = x.copy(), None, 0, 0
        for _ in range(400):
            x, lp, c = sampler.slice_step(x, f, 1.0, r1, left=left)
            xb, lpb, b, picks = sampler.slice_step_batched(xb, lpb, f, 1.0, r2, left=left)
            assert np.array_equal(x, xb) and np.array_equal(lp, lpb)
            assert (picks[1] >= -1).all()
            calls += c
            batches += b
        assert r1.random() == r2.random()
        assert batches < calls / 3


def test_batched_sweep_equals_the_sequential_one():
    S = 5

    def loglik(set_node, params):
        v, sv, a, sz, t, st = params[0], params[1], params[2], params[4], params[5], params[6]
        return -((v - 1.0) ** 2 + (a - 1.2) ** 2 * 4 + (t - 0.3) ** 2 * 50 + (sv - 1) ** 2 + (sz - .2) ** 2 + (st - .1) ** 2) + 0.01 * set_node

    for C in (1, 3):
        m1 = sampler.HierarchicalDDM(S, loglik, chains=C, seed=7, batched=False)
        m2 = sampler.HierarchicalDDM(S, loglik, chains=C, seed=7, batched=True)
        tr1, _ = m1.sample(25)
        tr2, _ = m2.sample(25)
        assert all(np.array_equal(tr1[k], tr2[k]) for k in tr1)
        assert m2.loglik_calls < m1.loglik_calls / 3


def test_likelihood_glue_pickles_without_device_handles():
    from hddm_b200 import likelihoods as L
    nodes = L.BatchedNodes([np.array([0.5, -0.7]), np.array([0.9])])
    assert nodes.node_off.tolist() == [0, 2, 3]
    nodes._trials = object()  # pretend a device handle exists
    clone = pickle.loads(pickle.dumps(nodes))
    assert clone._trials is None and np.array_equal(clone._x, nodes._x)
    cache = pickle.loads(pickle.dumps(L.DeviceTrialCache(8)))
    assert cache._handles == {} and cache.max_entries == 8
    p = nodes.pack(v=[1.0, 2.0], sv=0.5, a=1.5, z=0.5, sz=0.1, t=0.3, st=0.1, p_outlier=0.05)
    assert p.shape == (9, 2) and p[8, 0] == 0.1 and p[0, 1] == 2.0
    W = L.generate_wfpt_stochastic_class()
    for name in ("pdf", "cdf", "cdf_vec", "random", "chisquare", "gsquare", "theoretical_quantiles",
                 "empirical_quantiles", "get_quantiles_stats", "set_quantiles_stats"):
        assert hasattr(W, name)
