"""Random-walk simulator on the GPU (hddm/generate.py:208-319).

Parity mode: with the caller's random numbers the kernel must reproduce the numpy restatement of the
reference loop -- same crossing step, RT equal to 1e-12 (every operation is un-contracted FP64 in the
reference's order; the only freedom left is the sqrt/division rounding, identical under IEEE).
Production mode (device Philox): reproducible per seed and distributed as the model's density, the same
check hddm/tests/test_generate.py:21-29 makes between the two samplers."""
import numpy as np
import pytest

from tests.util import draw_params

pytestmark = pytest.mark.gpu
NAMES = ("v", "sv", "a", "z", "sz", "t", "st")


@pytest.fixture(scope="module")
def O():
    from oracle import oracle
    oracle.build()
    return oracle


def test_supplied_randomness_reproduces_the_reference_loop(gpu):
    from hddm_b200 import posterior
    from oracle import sim_oracle
    rng = np.random.default_rng(61)
    D, S, K = 5, 24, 12000
    params = draw_params(rng, D)
    params[2] = rng.uniform(0.6, 1.4, D)   # keep walks short enough for K steps at dt = 1e-3
    params[1, 0] = 0.0                     # sv == 0: drift deviate unused
    params[4, 1] = 0.0
    params[6, 2] = 0.0
    u_init = np.concatenate([rng.random((D, S, 2)), rng.standard_normal((D, S, 1))], axis=2)
    u_steps = rng.random((D, S, K))
    got = posterior.simulate_drift_batch(params, S, dt=1e-3, u_init=u_init, u_steps=u_steps)
    for d in range(D):
        want = sim_oracle.gen_rts_from_simulated_drift(dict(zip(NAMES, params[:, d])), u_init[d], u_steps[d], dt=1e-3)
        assert np.array_equal(np.isnan(got[d]), np.isnan(want))
        m = ~np.isnan(want)
        assert m.sum() >= S - 2
        np.testing.assert_allclose(got[d][m], want[m], rtol=1e-12, atol=0)
        assert np.array_equal(np.sign(got[d][m]), np.sign(want[m]))


def test_reference_runs_reproduced(gpu, golden_dir):
    """The kernel against the reference's OWN outputs: tests/golden/ref_simdrift.npz holds ten runs of
    _gen_rts_from_simulated_drift (hddm/generate.py:208-319, four with v_switch) together with the random
    numbers each run consumed.  Same crossing boundary on every trial, RT to 1e-12."""
    from hddm_b200 import posterior
    from tests.test_oracle import simdrift_cases
    n = 0
    for k, p, S, dt, intra_sv, u_init, u_steps, want, _ in simdrift_cases(golden_dir):
        col = np.array([[p['v']], [p.get('sv', 0.)], [p['a']], [p['z']], [p.get('sz', 0.)], [p['t']], [p.get('st', 0.)]])
        switch = None
        if 'v_switch' in p:
            switch = np.array([[p['v_switch']], [p.get('V_switch', 0.)], [p['t_switch']]])
        init = u_init[None, :, :4 if switch is not None else 3]
        got = posterior.simulate_drift_batch(col, S, dt=dt, intra_sv=intra_sv, u_init=init, u_steps=u_steps[None],
                                             switch=switch)[0]
        assert not np.isnan(got).any(), k
        assert np.array_equal(np.sign(got), np.sign(want)), k
        np.testing.assert_allclose(got, want, rtol=1e-12, atol=0, err_msg="case %d" % k)
        n += 1
    assert n >= 10


def test_switch_variant_device_rng(gpu):
    """v_switch with the device RNG: a strong negative switched drift after 50 ms turns a mostly-upper
    process into a mostly-lower one; V_switch = 0 and t_switch beyond every walk leave the plain process."""
    from hddm_b200 import posterior
    p = dict(v=2.0, a=2.0, z=0.5, t=0.3, sv=0.2, sz=0.1, st=0.1)
    plain = posterior.gen_rts_from_simulated_drift(p, samples=4000, dt=1e-3, seed=9)
    late = posterior.gen_rts_from_simulated_drift(dict(p, v_switch=-4.0, t_switch=50.0), samples=4000, dt=1e-3, seed=9)
    assert np.array_equal(plain, late)
    flipped = posterior.gen_rts_from_simulated_drift(dict(p, v_switch=-4.0, V_switch=0.5, t_switch=0.05),
                                                     samples=4000, dt=1e-3, seed=9)
    assert (plain > 0).mean() > 0.9 and (flipped > 0).mean() < 0.35
    with pytest.raises(ValueError):
        posterior.gen_rts_from_simulated_drift(dict(p, v_switch=1.0, t_switch=1e-5), samples=5, dt=1e-3)


def test_walk_that_never_crosses_is_nan(gpu):
    from hddm_b200 import posterior
    params = np.array([[0.0], [0.0], [50.0], [0.5], [0.0], [0.3], [0.0]])
    r = posterior.simulate_drift_batch(params, 8, dt=1e-4, seed=1, max_steps=1000)
    assert np.isnan(r).all()


def test_device_rng_reproducible_and_distributed_as_the_density(gpu, O):
    from hddm_b200 import posterior
    rng = np.random.default_rng(62)
    D, S = 4, 12000
    params = draw_params(rng, D)
    params[1, 1] = 0.0
    params[4, 2] = 0.0
    params[6, 3] = 0.0
    a = posterior.simulate_drift_batch(params, S, dt=1e-4, seed=5)
    b = posterior.simulate_drift_batch(params, S, dt=1e-4, seed=5)
    c = posterior.simulate_drift_batch(params, S, dt=1e-4, seed=6)
    assert np.array_equal(a, b) and not np.array_equal(a, c)
    assert not np.isnan(a).any()
    grid = np.arange(-8.0, 8.0, 1e-3)
    for d in range(D):
        pdf = O.pdf_array(grid, *params[:, d], 1e-4, 0, 2, 2, 1, 1e-3, 0, 0)
        cdf = np.cumsum(pdf) * 1e-3
        xs = np.sort(a[d])
        F = np.interp(xs, grid, cdf)
        emp = np.arange(1, S + 1) / float(S)
        # KS at the 0.1 % level + discretisation of the walk (a +-sqrt(dt) lattice overshoots the boundary
        # by half a step on average: a bias of order sqrt(dt) = 1e-2 in space, ~1e-2 in probability)
        assert np.max(np.abs(emp - F)) < 1.95 / np.sqrt(S) + 1.5e-2, (d, np.max(np.abs(emp - F)))


def test_reference_shaped_single_call(gpu):
    from hddm_b200 import posterior
    p = dict(v=1.0, a=1.6, z=0.5, t=0.3, sv=0.5, sz=0.1, st=0.1)
    r = posterior.gen_rts_from_simulated_drift(p, samples=500, seed=3)
    assert r.shape == (500,) and (np.abs(r) > 0.3 - 0.05).all()
    assert 0.6 < (r > 0).mean() < 0.95
    sw = posterior.gen_rts_from_simulated_drift(dict(p, v_switch=1.0, t_switch=0.2), samples=5, seed=3)
    assert sw.shape == (5,) and not np.isnan(sw).any()


def test_stochastic_class_random_with_drift_sampling(gpu):
    """Wfpt.random() with sampling_method='drift' (hddm/likelihoods.py:76-81 -> hddm/generate.py:187-188),
    the path HDDMRegressor's posterior predictive uses (hddm/models/hddm_regression.py:62)."""
    from hddm_b200 import likelihoods as L
    cls = L.generate_wfpt_stochastic_class(sampling_method='drift', sampling_dt=1e-3)
    node = cls(value=np.zeros(400), v=1.5, sv=0.3, a=1.5, z=0.5, sz=0.1, t=0.3, st=0.1)
    np.random.seed(3)
    r = node.random()
    assert r.shape == (400,) and not np.isnan(r).any()
    assert (np.abs(r) >= 0.3 - 0.05 - 1e-9).all() and 0.7 < (r > 0).mean() < 0.99
    with pytest.raises(TypeError):
        L.generate_wfpt_stochastic_class(sampling_method='nope')(value=np.zeros(3), v=1, sv=0, a=1, z=.5, sz=0, t=.3, st=0).random()
