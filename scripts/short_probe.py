"""Short-segment throughput probe: T trials x P sets through wiener_like_batch (end to end, best of 5)."""
import sys, time
import numpy as np
sys.path.insert(0, '.')
from hddm_b200 import wfpt
from bench_paths import draw, rts, HDDM
rng = np.random.default_rng(1)
for T, P in ((1000, 10000), (600, 10000), (300, 20000), (1000, 3000)):
    x = rts(rng, T)
    params = np.concatenate([draw(rng, P), np.full((1, P), 0.05), np.full((1, P), 0.1)])
    wfpt.wiener_like_batch(x, params, **HDDM)
    best = 1e9
    for _ in range(5):
        t0 = time.perf_counter(); wfpt.wiener_like_batch(x, params, **HDDM); best = min(best, time.perf_counter() - t0)
    print("%d x %d: %.4g evals/s (%.3f ms)" % (T, P, T * P / best, best * 1e3))
