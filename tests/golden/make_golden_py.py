"""Fixtures from the reference's own PYTHON functions on the path (run in the build container):

  * ref_ppc_stats.npz -- `gen_ppc_stats()` of /root/reference/hddm/utils.py:241-267, its fifteen lambdas
    applied to seeded RT sets (both boundaries, one boundary empty, ties, zeros, a single RT);
  * ref_simdrift.npz  -- `_gen_rts_from_simulated_drift` of /root/reference/hddm/generate.py:208-319 run with
    numpy's global RNG seeded, incl. the `v_switch` / `V_switch` variant (:234-241,272-276,283-285); the
    random numbers it consumed are stored beside its outputs (recorded at `numpy.random.rand`, replayed
    from the seed for the scipy `uniform.rvs` / `norm.rvs` draws and cross-checked against the recording).

hddm/utils.py and hddm/generate.py import kabuki / pymc / matplotlib / hddm at module top only; neither
function touches them, so the two files are executed UNMODIFIED, by path, with empty stand-in modules
for those names in sys.modules.  Nothing of the reference is copied.  /root/reference does not exist
on the GPU box; the .npz files do.          python tests/golden/make_golden_py.py
"""
import importlib.util
import os
import sys
import types
import warnings

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = "/root/reference/hddm"


def load_reference_module(name):
    """Execute /root/reference/hddm/<name>.py as a stand-alone module."""
    stubs = {}
    for mod in ("kabuki", "kabuki.analyze", "pymc", "matplotlib", "matplotlib.pyplot", "hddm"):
        if mod not in sys.modules:
            stubs[mod] = types.ModuleType(mod)
    if "kabuki.analyze" in stubs:
        stubs["kabuki.analyze"].post_pred_gen = stubs["kabuki.analyze"].post_pred_compare_stats = None
    if "kabuki" in stubs and "kabuki.analyze" in stubs:
        stubs["kabuki"].analyze = stubs["kabuki.analyze"]
    if "matplotlib" in stubs and "matplotlib.pyplot" in stubs:
        stubs["matplotlib"].pyplot = stubs["matplotlib.pyplot"]
    sys.modules.update(stubs)
    try:
        spec = importlib.util.spec_from_file_location("_hddm_ref_" + name, os.path.join(REF, name + ".py"))
        mod = importlib.util.module_from_spec(spec)
        spec.loader.exec_module(mod)
    finally:
        for k in stubs:
            sys.modules.pop(k, None)
    return mod


# ---------------------------------------------------------------------------------------------------------
def ppc_cases():
    rng = np.random.default_rng(8101)
    cases = []
    for n in (1, 2, 3, 7, 50, 200, 201, 1000, 4096):
        t = 0.25 + rng.gamma(2.0, 0.25, n)
        cases.append(t * np.where(rng.random(n) < 0.7, 1.0, -1.0))
    cases.append(0.3 + rng.gamma(2.0, 0.25, 120))                    # lower boundary empty
    cases.append(-(0.3 + rng.gamma(2.0, 0.25, 120)))                 # upper boundary empty
    tie = np.round(0.3 + rng.gamma(2.0, 0.25, 300), 1) * np.where(rng.random(300) < 0.5, 1.0, -1.0)
    cases.append(tie)                                                # heavy ties (0.1 s grid)
    z = 0.3 + rng.gamma(2.0, 0.25, 64)
    z[::8] = 0.0                                                     # zeros belong to neither boundary
    z[1::8] *= -1
    cases.append(z)
    cases.append(np.array([0.5]))
    cases.append(np.array([-0.5]))
    cases.append(np.array([0.7, -0.4]))
    cases.append(np.full(33, 0.61))                                  # zero variance
    # simulated data sets as gen_rts_from_cdf grids give them (multiples of dt plus a delay)
    for n in (200, 500):
        g = np.round(rng.gamma(2.0, 0.25, n), 4) + 0.2871
        cases.append(g * np.where(rng.random(n) < 0.85, 1.0, -1.0))
    return cases


def make_ppc():
    U = load_reference_module("utils")
    stats = U.gen_ppc_stats()
    names = list(stats.keys())
    out = {"names": np.array(names), "n": np.array(0)}
    cases = ppc_cases()
    with warnings.catch_warnings():
        warnings.simplefilter("ignore")   # np.mean / np.std of an empty selection warn and give nan
        for k, x in enumerate(cases):
            out["x_%d" % k] = x
            out["y_%d" % k] = np.array([float(f(x)) for f in stats.values()], dtype=np.float64)
    out["n"] = np.array(len(cases))
    np.savez_compressed(os.path.join(HERE, "ref_ppc_stats.npz"), **out)
    return len(cases), names


# ---------------------------------------------------------------------------------------------------------
class RandRecorder(object):
    """Stands in for numpy.random.rand while the reference function runs: same draws, kept."""

    def __init__(self):
        self.real = np.random.rand
        self.blocks = []

    def __call__(self, *shape):
        r = self.real(*shape)
        self.blocks.append(r.copy())
        return r


def run_simdrift(G, params, samples, dt, intra_sv, seed):
    """One reference call; returns rts and the random numbers it consumed, per sample:
    u_init [samples, 4] = delay uniform, start-point uniform, standard-normal drift deviate, standard-normal
    deviate of the switched drift (0 where the reference drew none), u_steps [samples, max_steps] (walk
    uniforms in order, NaN-padded), n_blocks [samples]."""
    rec = RandRecorder()
    np.random.seed(seed)
    np.random.rand = rec
    try:
        rts, drifts = G._gen_rts_from_simulated_drift(dict(params), samples=samples, dt=dt, intra_sv=intra_sv)
    finally:
        np.random.rand = rec.real
    nn = 1000
    if 'v_switch' in params:
        nn = int(round(params['t_switch'] / dt))          # hddm/generate.py:236-239
    # blocks per sample: the reference keeps drawing nn steps until a block holds a crossing; recover the
    # count from the crossing time it interpolated (:304-315) -- cross-checked below against a replay
    u_init = np.zeros((samples, 4))
    # replay the scipy draws from the seed (uniform.rvs -> RandomState.uniform(0, 1, n), norm.rvs ->
    # RandomState.standard_normal(); both on numpy's global state, in the reference's order :244-277)
    np.random.seed(seed)
    if 'st' in params:
        u_init[:, 0] = np.random.uniform(0.0, 1.0, samples)
    if 'sz' in params:
        u_init[:, 1] = np.random.uniform(0.0, 1.0, samples)
    blocks = list(rec.blocks)
    per_sample = []
    k = 0
    a = params['a']
    for i in range(samples):
        if 'sv' in params and params['sv'] != 0:
            u_init[i, 2] = np.random.standard_normal()
        if 'v_switch' in params and 'V_switch' in params and params['V_switch'] != 0:
            u_init[i, 3] = np.random.standard_normal()
        # how many blocks did sample i use?  |rt_i| - delay_i = crossing time in [ (it-1) nn dt, it nn dt ]
        st = params.get('st', 0.) if 'st' in params else 0.
        delay = (params['t'] + st * u_init[i, 0]) - st / 2. if 'st' in params else params['t']
        # and the crossing step (0-based, global) is ceil(crossing time / dt): the interpolation lands inside it
        cross_t = abs(rts[i]) - delay
        it = max(0, int(np.ceil(cross_t / dt - 1e-6))) // nn + 1
        mine = []
        for _ in range(it):
            again = np.random.rand(nn)                    # keeps the global state in step with the reference
            assert np.array_equal(again, blocks[k]), "replay diverged from the recording"
            mine.append(blocks[k])
            k += 1
        per_sample.append(np.concatenate(mine))
    assert k == len(blocks), (k, len(blocks))
    max_steps = max(len(u) for u in per_sample)
    u_steps = np.full((samples, max_steps), np.nan)
    for i, u in enumerate(per_sample):
        u_steps[i, :len(u)] = u
    n_blocks = np.array([len(u) // nn for u in per_sample])
    return np.asarray(rts, dtype=np.float64), u_init, u_steps, n_blocks


def make_simdrift():
    G = load_reference_module("generate")
    cases = [
        # (params, samples, dt, intra_sv, seed)
        (dict(v=1.2, a=1.4, z=0.5, t=0.3), 24, 1e-3, 1.0, 11),                                    # no sv/sz/st keys
        (dict(v=0.8, a=1.2, z=0.45, t=0.35, sv=0.6, sz=0.2, st=0.15), 32, 1e-3, 1.0, 12),          # full DDM
        (dict(v=-1.5, a=1.0, z=0.55, t=0.25, sv=0.0, sz=0.1, st=0.1), 24, 1e-3, 1.0, 13),          # sv key, == 0
        (dict(v=0.3, a=2.0, z=0.5, t=0.4, sv=1.0, sz=0.3, st=0.2), 16, 1e-3, 1.0, 14),             # long walks: > 1 block
        (dict(v=0.5, a=1.1, z=0.5, t=0.3, sv=0.4, st=0.1), 16, 1e-4, 1.0, 15),                     # dt = 1e-4, no sz key
        (dict(v=1.0, a=1.3, z=0.5, t=0.3, sv=0.3, sz=0.1, st=0.1), 16, 1e-3, 0.7, 16),             # intra_sv != 1
        (dict(v=1.5, a=1.6, z=0.5, t=0.3, v_switch=-2.0, t_switch=0.2), 24, 1e-3, 1.0, 21),        # switch, fixed
        (dict(v=1.5, a=1.6, z=0.5, t=0.3, sv=0.5, sz=0.1, st=0.1, v_switch=-2.5, V_switch=0.8, t_switch=0.15),
         32, 1e-3, 1.0, 22),                                                                       # switch, all noise
        (dict(v=0.2, a=1.8, z=0.4, t=0.3, sv=0.3, v_switch=3.0, V_switch=0.0, t_switch=0.0674), 24, 1e-3, 1.0, 23),
        (dict(v=-0.5, a=1.5, z=0.5, t=0.3, sz=0.2, v_switch=2.0, V_switch=0.5, t_switch=0.25), 12, 1e-4, 1.0, 24),
    ]
    keys = ("v", "sv", "a", "z", "sz", "t", "st", "v_switch", "V_switch", "t_switch")
    out = {"n": np.array(len(cases)), "keys": np.array(keys)}
    for k, (p, samples, dt, intra_sv, seed) in enumerate(cases):
        rts, u_init, u_steps, n_blocks = run_simdrift(G, p, samples, dt, intra_sv, seed)
        out["par_%d" % k] = np.array([p.get(name, np.nan) for name in keys])   # NaN: key absent
        out["cfg_%d" % k] = np.array([samples, dt, intra_sv, seed])
        out["rts_%d" % k] = rts
        out["u_init_%d" % k] = u_init
        out["u_steps_%d" % k] = u_steps
        out["blocks_%d" % k] = n_blocks
    np.savez_compressed(os.path.join(HERE, "ref_simdrift.npz"), **out)
    return len(cases)


if __name__ == "__main__":
    n, names = make_ppc()
    print("ref_ppc_stats.npz: %d data sets x %d statistics (%s ...)" % (n, len(names), ", ".join(names[:4])))
    print("ref_simdrift.npz: %d reference runs" % make_simdrift())
