"""Posterior-predictive batch (BASELINE config 5): many posterior draws per call.

CPU part: the statistics oracle (a one-for-one restatement of hddm/utils.py:241-267) on hand-checkable
data.  GPU part: the batched CUDA path against (a) D scalar calls of the already-pinned single-draw entry
points (bit for bit), (b) the C oracle of gen_rts_from_cdf / dmat_cdf_array on the same uniforms, (c) the
statistics oracle, and -- for the device-RNG variant, whose stream cannot match numpy's -- distributional
properties: reproducibility per seed, and agreement of the simulated quantiles with the analytic CDF."""
import numpy as np
import pytest

from tests.util import assert_parity, cdf_noise_band, draw_params


@pytest.fixture(scope="module")
def O():
    from oracle import oracle
    oracle.build()
    return oracle


def test_stats_oracle_hand_values():
    from oracle.ppc_oracle import ppc_stats
    x = np.array([0.5, 0.7, 0.9, 1.1, 1.3, -0.4, -0.6, -0.8, 0.0])
    s = ppc_stats(x)
    assert s[0] == 5 / 9.
    np.testing.assert_allclose(s[1], 0.9)
    np.testing.assert_allclose(s[2], np.sqrt(0.08))
    np.testing.assert_allclose(s[3:8], [0.58, 0.74, 0.9, 1.06, 1.22])   # idx = q/100*(5-1)
    np.testing.assert_allclose(s[8], -0.6)                                # mean of the SIGNED lower RTs
    np.testing.assert_allclose(s[10:15], [0.44, 0.52, 0.6, 0.68, 0.76])
    only_upper = ppc_stats(np.array([0.5, 0.6]))
    assert only_upper[0] == 1.0 and np.isnan(only_upper[8:]).all()


# ---- GPU ---------------------------------------------------------------------------------------

def _draws(rng, D):
    p = draw_params(rng, D)
    p[6, ::4] = 0.0   # some draws without st (no delay uniforms are consumed for them)
    p[4, 1::5] = 0.0  # and without sz
    p[1, 2::7] = 0.0  # and without sv
    return np.ascontiguousarray(p)


@pytest.mark.gpu
def test_batch_equals_scalar_calls_and_oracle(gpu, O):
    from hddm_b200 import posterior, wfpt
    rng = np.random.default_rng(41)
    D, S = 23, 300
    params = _draws(rng, D)
    f, u = rng.random((D, S)), rng.random((D, S))
    rts, stats = posterior.gen_rts_from_cdf_batch(params, S, -5.0, 5.0, 1e-3, f=f, u_delay=u, return_stats=True)
    exact = 0
    for d in range(D):
        one = wfpt.gen_rts_from_cdf(*params[:, d], S, -5.0, 5.0, 1e-3, f=f[d], u_delay=u[d])
        assert np.array_equal(one, rts[d]), d  # same kernels, same order: bit for bit
        want = O.gen_rts_from_cdf(*params[:, d], S, -5.0, 5.0, 1e-3, f=f[d], u_delay=u[d])
        close = np.abs(rts[d] - want) <= 1e-12
        exact += close.sum()
        assert np.max(np.abs(rts[d] - want)) <= 1e-3 + 1e-12  # at most one grid step (discrete lookup)
    assert exact >= 0.995 * D * S
    # numpy's global stream, consumed draw by draw as D reference calls would
    np.random.seed(7)
    a = posterior.gen_rts_from_cdf_batch(params, S, -5.0, 5.0, 1e-3)
    np.random.seed(7)
    b = np.stack([wfpt.gen_rts_from_cdf(*params[:, d], S, -5.0, 5.0, 1e-3) for d in range(D)])
    assert np.array_equal(a, b)


@pytest.mark.gpu
@pytest.mark.parametrize("S", [1, 2, 37, 256, 1000, 5000])
def test_stats_against_oracle(gpu, S):
    from hddm_b200 import posterior
    from oracle.ppc_oracle import ppc_stats
    rng = np.random.default_rng(100 + S)
    D = 12
    rts = (0.3 + rng.gamma(2.0, 0.25, (D, S))) * np.where(rng.random((D, S)) < 0.7, 1.0, -1.0)
    rts[0] = np.abs(rts[0])           # no lower-boundary responses -> NaN block
    rts[1] = -np.abs(rts[1])          # no upper-boundary responses
    if S > 2:
        rts[2, :2] = 0.0              # zeros belong to neither boundary but count in the accuracy
        rts[3, 1] = rts[3, 0]         # ties
    got = posterior.ppc_stats(rts)
    want = np.stack([ppc_stats(r) for r in rts])
    assert np.array_equal(np.isnan(got), np.isnan(want))
    m = ~np.isnan(want)
    # sums are reduced in a different (fixed) order than numpy's pairwise sum: 1e-12, not bitwise
    np.testing.assert_allclose(got[m], want[m], rtol=1e-12, atol=1e-15)
    assert np.array_equal(got[:, 0], want[:, 0])  # accuracy is a ratio of counts: exact


@pytest.mark.gpu
def test_stats_against_reference_fixture(gpu, golden_dir):
    """k_ppc_stats against the reference's OWN gen_ppc_stats() (hddm/utils.py:241-267) outputs,
    tests/golden/ref_ppc_stats.npz (made by tests/golden/make_golden_py.py from the unmodified file)."""
    import os
    from hddm_b200 import posterior
    d = np.load(os.path.join(golden_dir, "ref_ppc_stats.npz"))
    for k in range(int(d["n"])):
        got, want = posterior.ppc_stats(d["x_%d" % k])[0], d["y_%d" % k]
        assert np.array_equal(np.isnan(got), np.isnan(want)), k
        m = ~np.isnan(want)
        np.testing.assert_allclose(got[m], want[m], rtol=1e-12, atol=1e-15, err_msg="data set %d" % k)
        assert got[0] == want[0]


@pytest.mark.gpu
def test_fused_stats_equal_stats_of_returned_rts(gpu):
    from hddm_b200 import posterior
    rng = np.random.default_rng(3)
    D, S = 40, 200
    params = _draws(rng, D)
    f, u = rng.random((D, S)), rng.random((D, S))
    rts, stats = posterior.gen_rts_from_cdf_batch(params, S, -5.0, 5.0, 1e-3, f=f, u_delay=u, return_stats=True)
    again = posterior.ppc_stats(rts)
    assert np.array_equal(np.isnan(stats), np.isnan(again))
    assert np.array_equal(stats[~np.isnan(stats)], again[~np.isnan(again)])
    only = posterior.gen_rts_from_cdf_batch(params, S, -5.0, 5.0, 1e-3, f=f, u_delay=u, return_rts=False,
                                            return_stats=True)
    assert np.array_equal(only[~np.isnan(only)], stats[~np.isnan(stats)])


@pytest.mark.gpu
def test_device_rng_is_reproducible_and_follows_the_cdf(gpu, O):
    from hddm_b200 import posterior
    rng = np.random.default_rng(8)
    D, S = 6, 16000
    params = _draws(rng, D)
    a = posterior.gen_rts_from_cdf_batch(params, S, -6.0, 6.0, 1e-4, seed=123)
    b = posterior.gen_rts_from_cdf_batch(params, S, -6.0, 6.0, 1e-4, seed=123)
    c = posterior.gen_rts_from_cdf_batch(params, S, -6.0, 6.0, 1e-4, seed=124)
    assert np.array_equal(a, b) and not np.array_equal(a, c)
    # a sub-batch reproduces its rows only when offsets agree; the first draw always does
    first = posterior.gen_rts_from_cdf_batch(params[:, :1].copy(), S, -6.0, 6.0, 1e-4, seed=123)
    assert np.array_equal(first[0], a[0])
    grid = np.arange(-6.0, 6.0, 1e-3)
    for d in range(D):
        # Kolmogorov distance between the simulated signed RTs and the CDF obtained by integrating the
        # oracle's full_pdf (with t and st) on a fine grid.  (Not cdfdif: the reference's own 6-node
        # Gauss-Hermite CDF is off by 0.02 from its own density for sv ~ 2.5, e.g. draw 1 here.)
        pdf = O.pdf_array(grid, *params[:, d], 1e-4, 0, 2, 2, 1, 1e-3, 0, 0)
        cdf = np.cumsum(pdf) * 1e-3
        assert abs(cdf[-1] - 1) < 1e-3
        xs = np.sort(a[d])
        F = np.interp(xs, grid, cdf)
        emp = np.arange(1, S + 1) / float(S)
        # KS 0.1 % level + slack for the rectangle rules on both sides (dt * peak density)
        assert np.max(np.abs(emp - F)) < 1.95 / np.sqrt(S) + 4e-3, d


@pytest.mark.gpu
def test_cdf_batch_against_scalar_calls_and_oracle(gpu, O):
    from hddm_b200 import cdfdif_wrapper, posterior
    rng = np.random.default_rng(12)
    D, n = 29, 11
    par7 = draw_params(rng, D)
    par7[4] = np.maximum(par7[4], 0.02)  # full DDM: away from the 1e-10-epsilon regime (see cdf_noise_band)
    par7[6] = np.maximum(par7[6], 0.02)
    params = np.concatenate([par7, np.full((1, D), 0.05), np.full((1, D), 0.1)])
    params[3, 5] = 0.97                  # out of support: z + sz/2 > 1
    params[4, 5] = 0.2
    x = (par7[5] + par7[6] / 2 + rng.gamma(2.0, 0.3, (n, D))).T * np.where(rng.random((D, n)) < 0.5, 1.0, -1.0)
    x = np.ascontiguousarray(x)
    x[9, 3] = 7.0                        # beyond 1/(2 w_outlier) = 5: the reference asserts
    y, status = posterior.dmat_cdf_batch(x, params, return_status=True)
    for d in range(D):
        if d == 5:
            assert status[d] == 3 and np.isnan(y[d]).all()
            with pytest.raises(ValueError):
                cdfdif_wrapper.dmat_cdf_array(x[d], *params[:, d])
            continue
        if d == 9:
            assert status[d] == 4 and np.isnan(y[d]).all()
            with pytest.raises(AssertionError):
                cdfdif_wrapper.dmat_cdf_array(x[d], *params[:, d])
            continue
        assert status[d] == 0
        assert np.array_equal(y[d], cdfdif_wrapper.dmat_cdf_array(x[d], *params[:, d]))
        want = O.dmat_cdf_array(x[d], *params[:, d])
        band = cdf_noise_band(O, x[d], params[:7, d], params[7, d], params[8, d])
        assert (np.abs(y[d] - want) <= 1e-10 * np.abs(want) + 4 * band + 1e-14).all(), d


@pytest.mark.gpu
def test_quantile_objectives_match_the_stochastic_class(gpu):
    """chi-square / G-square for many parameter sets in one launch == the per-node methods of the wfpt
    stochastic class (hddm/likelihoods.py:198-239), which make one cdfdif call each."""
    from hddm_b200 import likelihoods as L, posterior, wfpt
    rng = np.random.default_rng(19)
    cls = L.generate_wfpt_stochastic_class()
    D = 7
    par7 = draw_params(rng, D)
    emp, freq, ns, chi, gsq = [], [], [], [], []
    for d in range(D):
        data = wfpt.gen_rts_from_cdf(*par7[:, d], 300, -5.0, 5.0, 1e-3, f=rng.random(300), u_delay=rng.random(300))
        other = par7[:, (d + 1) % D]   # evaluate the objective at a different parameter set
        node = cls(value=data, **dict(zip(("v", "sv", "a", "z", "sz", "t", "st"), other)), p_outlier=0.05)
        node.compute_quantiles_stats()
        emp.append(node._emp_rt); freq.append(node._freq_obs); ns.append(node._n_samples)
        chi.append(node.chisquare()); gsq.append(node.gsquare())
    params = np.concatenate([np.roll(par7, -1, axis=1), np.full((1, D), 0.05), np.full((1, D), 0.1)])
    c, g = posterior.quantile_objectives(np.stack(emp), np.stack(freq), np.array(ns), params)
    np.testing.assert_array_equal(c, np.array(chi))
    np.testing.assert_array_equal(g, np.array(gsq))


@pytest.mark.gpu
def test_device_rng_rows_do_not_depend_on_how_draws_are_split(gpu):
    """Sharding invariance on one GPU: the draws evaluated as one batch, or as two blocks that pass their
    global first_draw, give the same simulated data sets and statistics (what makes the multi-GPU table
    independent of the world size)."""
    from hddm_b200 import posterior
    rng = np.random.default_rng(21)
    D, S = 17, 150
    params = _draws(rng, D)
    rts, stats = posterior.gen_rts_from_cdf_batch(params, S, -5.0, 5.0, 1e-3, seed=9, return_stats=True)
    cut = 6
    a = posterior.gen_rts_from_cdf_batch(np.ascontiguousarray(params[:, :cut]), S, -5.0, 5.0, 1e-3, seed=9)
    b = posterior.gen_rts_from_cdf_batch(np.ascontiguousarray(params[:, cut:]), S, -5.0, 5.0, 1e-3, seed=9, first_draw=cut)
    assert np.array_equal(np.concatenate([a, b]), rts)
    table = posterior.post_pred_stats_sharded(params, S, -5.0, 5.0, 1e-3, seed=9)  # world size 1
    assert np.array_equal(np.isnan(table), np.isnan(stats))
    assert np.array_equal(table[~np.isnan(table)], stats[~np.isnan(stats)])


def _ppc_gloo_worker(rank, world, port, D, q):
    import os
    import sys
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from hddm_b200 import posterior
    table = posterior.post_pred_stats_sharded(_ppc_params(D), 50, evaluate=_ppc_standin)
    q.put((rank, table.copy()))
    dist.barrier()
    dist.destroy_process_group()


def _ppc_params(D):
    return np.ascontiguousarray(draw_params(np.random.default_rng(77), D))


def _ppc_standin(block, first_draw):
    """Stand-in for the CUDA launch (tests only): the statistics oracle on RTs keyed by the GLOBAL draw index."""
    from oracle.ppc_oracle import ppc_stats
    rows = []
    for d in range(block.shape[1]):
        r = np.random.default_rng(1000 + first_draw + d)
        rts = (block[5, d] + r.gamma(2.0, 0.25, 50)) * np.where(r.random(50) < 0.7, 1.0, -1.0)
        rows.append(ppc_stats(rts))
    return np.stack(rows)


@pytest.mark.parametrize("D", [9, 1, 6])
def test_sharded_post_pred_stats_over_gloo_world2(D):
    """N>1 host path of the posterior-predictive batch: draws split over 2 ranks, one all_gather of the
    [D, 15] statistics, every rank sees the single-process table bit for bit."""
    import socket
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_ppc_gloo_worker, args=(r, 2, port, D, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = _ppc_standin(_ppc_params(D), 0)
    for r in range(2):
        assert got[r].shape == (D, 15)
        assert np.array_equal(np.isnan(got[r]), np.isnan(want))
        assert np.array_equal(got[r][~np.isnan(want)], want[~np.isnan(want)])


def _cdf_standin(x_block, params_block):
    """Stand-in for the CUDA launch (tests only): a cheap function of each row's own inputs."""
    return np.tanh(x_block * params_block[0][:, None]) + params_block[2][:, None]


def _cdf_gloo_worker(rank, world, port, D, q):
    import os
    import sys
    import torch.distributed as dist
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
    from hddm_b200 import posterior
    x, params = _cdf_inputs(D)
    q.put((rank, posterior.dmat_cdf_sharded(x, params, evaluate=_cdf_standin).copy()))
    dist.barrier()
    dist.destroy_process_group()


def _cdf_inputs(D):
    rng = np.random.default_rng(78)
    par = np.concatenate([draw_params(rng, D), np.full((1, D), 0.05), np.full((1, D), 0.1)])
    return rng.uniform(-3, 3, (D, 11)), np.ascontiguousarray(par)


@pytest.mark.parametrize("D", [7, 1, 8])
def test_sharded_cdf_quantiles_over_gloo_world2(D):
    """N>1 host path of the cdfdif quantile batch: parameter sets split over 2 ranks (uneven, a rank with nothing,
    even), one all_gather of the [D, n] probabilities, every rank sees the single-process table bit for bit."""
    import socket
    import torch.multiprocessing as mp
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_cdf_gloo_worker, args=(r, 2, port, D, q)) for r in range(2)]
    for p in procs:
        p.start()
    got = dict(q.get(timeout=120) for _ in range(2))
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    want = _cdf_standin(*_cdf_inputs(D))
    for r in range(2):
        assert got[r].shape == (D, 11)
        assert np.array_equal(got[r], want)


@pytest.mark.gpu
def test_sharded_cdf_single_process_equals_batch(gpu):
    from hddm_b200 import posterior
    x, params = _cdf_inputs(40)
    params[4] = np.maximum(params[4], 0.02)
    params[6] = np.maximum(params[6], 0.02)
    x = np.clip(x, -4.9, 4.9)
    a, b = posterior.dmat_cdf_sharded(x, params), posterior.dmat_cdf_batch(x, params)
    assert np.array_equal(np.isnan(a), np.isnan(b)) and np.array_equal(a[~np.isnan(a)], b[~np.isnan(b)])


@pytest.mark.gpu
def test_batch_with_simple_and_single_time_draws(gpu, O):
    """Posterior draws without sz (at or below the 1e-3 snap) take the one-pdf_sv path, draws with sz the single-node
    path: the batch still equals one scalar call per draw bit for bit, and the reference's lookup on >= 99.5 %."""
    from hddm_b200 import posterior, wfpt
    rng = np.random.default_rng(43)
    D, S = 12, 400
    params = _draws(rng, D)
    params[4, 0::3] = 0.0
    params[4, 1::3] = 5e-4
    params[1, 0::4] = 0.0          # sv == 0 as well
    f, u = rng.random((D, S)), rng.random((D, S))
    rts = posterior.gen_rts_from_cdf_batch(params, S, -5.0, 5.0, 1e-3, f=f, u_delay=u)
    exact = 0
    for d in range(D):
        one = wfpt.gen_rts_from_cdf(*params[:, d], S, -5.0, 5.0, 1e-3, f=f[d], u_delay=u[d])
        assert np.array_equal(one, rts[d]), d
        want = O.gen_rts_from_cdf(*params[:, d], S, -5.0, 5.0, 1e-3, f=f[d], u_delay=u[d])
        exact += (np.abs(rts[d] - want) <= 1e-12).sum()
        assert np.max(np.abs(rts[d] - want)) <= 1e-3 + 1e-12
    assert exact >= 0.995 * D * S


@pytest.mark.gpu
def test_density_table_points_make_the_reference_decisions(gpu, O):
    """The points of a gen_rts_from_cdf table (t = st = 0: one decision time each) go through the lean single-node
    evaluator: densities at 1e-10 and, through the device-pointer entry point's counter, exactly the pdf_sv
    evaluations the reference's recursion makes (first level, refinements, the points around zero)."""
    import ctypes as C
    import torch
    from hddm_b200 import _lib
    from tests.util import RTOL_TRIAL, assert_parity
    L = _lib.lib()
    _lib.check(L.hddm_b200_init(0))
    dev = torch.device("cuda", 0)
    x = np.arange(-4, 4, 1e-3)
    xd = torch.from_numpy(x).to(dev)
    rng = np.random.default_rng(31)
    for k in range(6):
        v, sv, a, z = rng.uniform(-4, 4), rng.uniform(0, 2.5), rng.uniform(0.5, 2.0), rng.uniform(0.4, 0.6)
        sz = (0.0, 0.05, 0.2, 0.39, 0.3, 0.1)[k]
        y = torch.empty_like(xd)
        ev = torch.zeros(1, dtype=torch.int64, device=dev)
        _lib.check(L.hddm_pdf_array_dev(C.c_void_p(xd.data_ptr()), x.size, v, sv, a, z, sz, 0.0, 0.0, 1e-4, 0, 2, 2, 1, 1e-3,
                                        0.0, 0.0, C.c_void_p(y.data_ptr()), C.c_void_p(ev.data_ptr()),
                                        C.c_void_p(torch.cuda.current_stream(dev).cuda_stream)))
        torch.cuda.synchronize(dev)
        O.pdfsv_calls(reset=True)
        want = O.pdf_array(x, v, sv, a, z, sz, 0.0, 0.0, err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3)
        calls = O.pdfsv_calls(reset=True)
        assert_parity(y.cpu().numpy(), want, RTOL_TRIAL, what="table set %d" % k)
        assert int(ev.item()) == calls, (k, int(ev.item()), calls)


@pytest.mark.gpu
def test_batch_edge_cases(gpu):
    """Empty and maximum sizes of the batched entry points (the reference's own edge behaviour: an empty
    trial vector gives 0.0, src/wfpt.pyx:62-75; empty batches give empty results)."""
    from hddm_b200 import posterior, wfpt
    rng = np.random.default_rng(2)
    none = np.zeros((7, 0))
    assert posterior.gen_rts_from_cdf_batch(none, 10, -3, 3, 1e-2, seed=1).shape == (0, 10)
    assert posterior.ppc_stats(np.zeros((0, 5))).shape == (0, 15)
    assert posterior.dmat_cdf_batch(np.zeros((0, 4)), np.zeros((9, 0))).shape == (0, 4)
    assert posterior.simulate_drift_batch(none, 5).shape == (0, 5)
    # the largest data set the statistics kernel takes, and one beyond
    big = (0.3 + rng.gamma(2.0, 0.25, (2, posterior.MAX_SAMPLES))) * np.where(rng.random((2, posterior.MAX_SAMPLES)) < 0.6, 1., -1.)
    from oracle.ppc_oracle import ppc_stats
    np.testing.assert_allclose(posterior.ppc_stats(big), np.stack([ppc_stats(r) for r in big]), rtol=1e-12)
    with pytest.raises(ValueError):
        posterior.ppc_stats(np.ones((1, posterior.MAX_SAMPLES + 1)))
    # batched likelihoods: no sets, no trials
    x = (0.8 + rng.gamma(2.0, 0.25, 50)) * np.where(rng.random(50) < 0.6, 1., -1.)
    assert wfpt.wiener_like_batch(x, np.zeros((9, 0))).shape == (0,)
    p = np.array([[1.], [.5], [2.], [.5], [.1], [.3], [.1], [.05], [.1]])
    assert wfpt.wiener_like_batch(np.zeros(0), p)[0] == 0.0
    assert wfpt.wiener_like_regression_batch(np.zeros(0), np.zeros((2, 0)), np.ones((2, 1)), p)[0] == 0.0
    one = wfpt.wiener_like_regression_batch(x, np.ones((1, 50)), np.array([[1.0]]), p, n_st=2, n_sz=2)
    assert one[0] == wfpt.wiener_like_multi(x, np.full(50, 1.0), .5, 2., .5, .1, .3, .1, 1e-4, ['v'], 2, 2, 1, 1e-3, .05, .1)
    l = np.zeros(0, dtype=np.int64)
    assert wfpt.wiener_like_rlddm_batch(np.zeros(0), l, np.zeros(0), l, np.zeros((12, 0))).shape == (0,)
    assert wfpt.wiener_like_rl_batch(l, np.zeros(0), l, np.zeros((7, 0))).shape == (0,)
