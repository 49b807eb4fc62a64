#!/usr/bin/env python
"""bench_ppc.py -- BASELINE.json configs[4]: posterior predictive for D posterior draws across N GPUs.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 bench_ppc.py [--draws D]

Strong scaling: the D draws (default 10000, x 200 simulated trials each, CDF table dt = 1e-4 on [-6, 6) as
Wfpt.random uses, hddm/likelihoods.py:76-81) are sharded in contiguous blocks over the ranks; each rank
builds its density tables, samples and reduces to the 15 gen_ppc_stats statistics on its GPU; one NCCL
all_gather moves 120 bytes per draw.  Timed end to end through hddm_b200.posterior.post_pred_stats_sharded
(host parameters in, host table out), wall clock between barriers, max over ranks.  One JSON line (rank 0).
"""
import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--draws", type=int, default=10000)
    ap.add_argument("--samples", type=int, default=200)
    ap.add_argument("--steps", type=int, default=3)
    args = ap.parse_args()
    import torch
    import torch.distributed as dist
    from hddm_b200 import _lib, posterior
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    torch.cuda.set_device(local)
    if world > 1:
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    _lib.check(_lib.lib().hddm_b200_init(local))
    rng = np.random.default_rng(20261104)
    D = args.draws
    params = np.stack([rng.uniform(-4, 4, D), rng.uniform(0, 2.5, D), rng.uniform(0.5, 2.0, D), rng.uniform(0.4, 0.6, D),
                       rng.uniform(0, 0.4, D), rng.uniform(0.2, 0.5, D), rng.uniform(0, 0.35, D)])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    table = posterior.post_pred_stats_sharded(params, args.samples, -6, 6, 1e-4, seed=1)  # warm-up
    times = []
    for _ in range(args.steps):
        barrier()
        t0 = time.perf_counter()
        table = posterior.post_pred_stats_sharded(params, args.samples, -6, 6, 1e-4, seed=1)
        barrier()
        times.append(time.perf_counter() - t0)
    t = torch.tensor([min(times)], dtype=torch.float64, device="cuda")
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    if rank == 0:
        import hashlib
        print(json.dumps({"metric": "posterior predictive draws/sec (gen_rts_from_cdf + gen_ppc_stats per draw)",
                          "value": D / float(t.item()), "unit": "draws/s", "n_gpus": world, "scaling": "strong",
                          "ms": 1e3 * float(t.item()), "dtype": "f64", "data": "synthetic",
                          "config": {"workload": "C5: %d posterior draws x %d trials, cdf table dt=1e-4 on [-6,6), "
                                                 "15 statistics per draw" % (D, args.samples)},
                          "table_sha1": hashlib.sha1(np.nan_to_num(table).tobytes()).hexdigest()}))
    if world > 1:
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
