import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from hddm_b200 import likelihoods as L, sampler
rts = bench.mcmc_data()
nodes = L.BatchedNodes(rts)
m = sampler.HierarchicalDDM(len(rts), nodes.logp, chains=16, seed=1)
m.sample_native(nodes, 40, trace=False)
_, dt = m.sample_native(nodes, 6, trace=False)
print(6 / dt)
