"""A short native-sampler run (profiling / latency probe): 1 chain and 16 chains of the C2 model."""
import sys, os, time
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from hddm_b200 import likelihoods as L, sampler
rts = bench.mcmc_data()
nodes = L.BatchedNodes(rts)
for chains, iters in ((1, int(sys.argv[1]) if len(sys.argv) > 1 else 200), (16, 30)):
    m = sampler.HierarchicalDDM(len(rts), nodes.logp, chains=chains, seed=1)
    m.sample_native(nodes, 5, trace=False)
    c0 = m.loglik_calls
    _, dt = m.sample_native(nodes, iters, trace=False)
    print(chains, "chains:", iters / dt, "it/s per chain,", (m.loglik_calls - c0) / iters, "launches/iter,", 1e6 * dt / (m.loglik_calls - c0), "us per launch")
