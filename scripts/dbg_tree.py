"""Debug build only (-DHDDM_DEBUG_TREE): why deferred trials leave the lean tree (sweep S-B)."""
import ctypes as C, sys, os
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import bench
from hddm_b200 import _lib, wfpt
L = C.CDLL(os.environ["HDDM_B200_LIB"])
x, params = bench.sweep_inputs(20000, 2000, 0.30, 20261104)
out = (C.c_ulonglong * 16)()
lp, ev = wfpt.wiener_like_batch(x, params, return_evals=True, **bench.HDDM)
L.hddm_debug_counters(out, 1)
lp, ev = wfpt.wiener_like_batch(x, params, return_evals=True, **bench.HDDM)
L.hddm_debug_counters(out, 1)
names = ["enter", "fail", "onset/NaN", "small<lo", "large>hi", "inner-bad large", "inner-bad small", "nodes", "neg", "refine"]
print({n: int(out[i]) for i, n in enumerate(names)}, "evals", ev, "trials", x.size * params.shape[1])
