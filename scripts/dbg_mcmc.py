"""Debug build only (-DHDDM_DEBUG_TREE): where the node-batch kernel spends MCMC launches."""
import ctypes as C, sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from hddm_b200 import likelihoods as L, sampler
lib = C.CDLL(os.environ["HDDM_B200_LIB"])
rts = bench.mcmc_data()
nodes = L.BatchedNodes(rts)
m = sampler.HierarchicalDDM(len(rts), nodes.logp, chains=16, seed=1)
out = (C.c_ulonglong * 16)()
names = ["tree enter", "tree fail", "onset/NaN", "small<lo", "large>hi", "inner-bad large", "inner-bad small", "nodes", "neg", "refine",
         "sets", "sets !lean", "!support", "!G.ok", "st/sz==0", "queued"]
for phase, n in (("burn-in 0-40", 40), ("40-60", 20)):
    lib.hddm_debug_counters(out, 1)
    m.sample_native(nodes, n, trace=False)
    lib.hddm_debug_counters(out, 1)
    print(phase, {k: int(out[i]) for i, k in enumerate(names)})
print({k: float(v.mean()) for k, v in m.g.items()})
