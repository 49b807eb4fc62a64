import time, numpy as np, sys
import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from hddm_b200 import likelihoods as L
import bench_mcmc
rts = bench_mcmc.make_data(20,200,1)
nodes = L.BatchedNodes(rts)
for C in (1,16,256):
    P=20*C
    set_node=np.tile(np.arange(20),C)
    params=nodes.pack(v=np.full(P,1.0),sv=0.8,a=1.5,z=0.5,sz=0.2,t=0.3,st=0.15,p_outlier=0.05)
    nodes.logp(set_node,params)
    t0=time.perf_counter()
    for _ in range(200): nodes.logp(set_node,params)
    dt=(time.perf_counter()-t0)/200
    print("C",C,"per call us",dt*1e6,"evals/s",P*200/dt)

# split: raw C-ABI call vs Python wrapper
import ctypes as C
from hddm_b200 import _lib
lib = _lib.lib()
h = nodes._handle()
for Cn in (1, 256):
    P = 20 * Cn
    set_node = np.tile(np.arange(20), Cn).astype(np.int64)
    params = np.ascontiguousarray(nodes.pack(v=np.full(P, 1.0), sv=0.8, a=1.5, z=0.5, sz=0.2, t=0.3, st=0.15, p_outlier=0.05))
    out = np.empty(P); ev = C.c_ulonglong(0)
    args = (h._h, nodes.node_off.ctypes.data_as(_lib.c_lp), 20, set_node.ctypes.data_as(_lib.c_lp), _lib.dptr(params), P,
            1e-4, 2, 2, 1, 1e-3, _lib.dptr(out), C.byref(ev))
    lib.hddm_trials_wiener_like_nodes(*args)
    t0 = time.perf_counter()
    for _ in range(300): lib.hddm_trials_wiener_like_nodes(*args)
    print("raw ABI call, C =", Cn, (time.perf_counter() - t0) / 300 * 1e6, "us")
