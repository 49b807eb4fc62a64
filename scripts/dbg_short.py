import sys, numpy as np
sys.path.insert(0, '.')
from tests.util import draw_rts, draw_params
from hddm_b200 import wfpt as W
HDDM = dict(err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3)
P = 1500
rng = np.random.default_rng(9500 + P)
lens = [0, 37, 200, 64, 311, 1000, 1]
xs = [draw_rts(rng, n, 0.40) for n in lens]
x = np.concatenate(xs)
node_off = np.concatenate([[0], np.cumsum(lens)])
set_node = rng.integers(0, len(lens), P)
par = draw_params(rng, P)
params = np.concatenate([par, np.full((1, P), 0.05), np.full((1, P), 0.1)])
params[7, ::7] = 0.0; params[4, ::11] = 0.0; params[6, ::13] = 0.0
h = W.Trials(x)
got, evals = h.wiener_like_nodes(node_off, set_node, params, return_evals=True, **HDDM)
got2, evals2 = h.wiener_like_nodes(node_off, set_node, params, return_evals=True, **HDDM)
print("repeat equal", np.array_equal(got, got2, equal_nan=True), evals, evals2)
small = []; se = 0
for lo in range(0, P, 500):
    g, e = h.wiener_like_nodes(node_off, set_node[lo:lo+500], np.ascontiguousarray(params[:, lo:lo+500]), return_evals=True, **HDDM)
    small.append(g); se += e
small = np.concatenate(small)
bad = np.where(~((got == small) | (np.isnan(got) & np.isnan(small))))[0]
print("differing sets", len(bad), "evals", evals, se)
for p in bad[:20]:
    print(p, "node", set_node[p], "len", lens[set_node[p]], got[p], small[p], (got[p]-small[p])/abs(small[p]), "sz", params[4,p], "st", params[6,p], "po", params[7,p])
