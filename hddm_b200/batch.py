"""Batched / sharded evaluation of the Wiener likelihood on device-resident data.

One process per GPU.  PyTorch is plumbing only: it owns the device buffers, the stream the
kernels are enqueued on and the `torch.distributed` process group.  The arithmetic happens in
libhddm_b200.so through the `_dev` entry points of include/hddm_b200.h.

Sharding: parameter sets (chains, subjects, proposals) are split into
contiguous blocks over ranks; every rank holds the trials its sets need; each per-set sum is
computed entirely on one GPU in a fixed order, so the gathered vector is bit-identical for any
world size.  The only collective is one all_gather of the per-set log-likelihoods.
"""
import ctypes as C

import numpy as np

from . import _lib

HDDM_DEFAULTS = dict(err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3)


def shard_bounds(n_sets, world_size, rank):
    """Contiguous block [lo, hi) of `n_sets` owned by `rank` (first n_sets % world ranks get one more)."""
    base, extra = divmod(int(n_sets), int(world_size))
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard_sizes(n_sets, world_size):
    return [shard_bounds(n_sets, world_size, r)[1] - shard_bounds(n_sets, world_size, r)[0]
            for r in range(world_size)]


class DeviceSweep(object):
    """Trials resident in HBM + reusable scratch for repeated batched launches on one GPU."""

    def __init__(self, x, device=None, seg=None, sort=True):
        import torch
        self.torch = torch
        if not torch.cuda.is_available():
            raise RuntimeError("hddm_b200.batch needs a CUDA device; there is no CPU fallback")
        self.device = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
        x = _lib.as_rt_buffer(x) if isinstance(x, np.ndarray) else x
        self.x = (torch.from_numpy(x) if isinstance(x, np.ndarray) else x).to(self.device, torch.float64).contiguous()
        self.n = int(self.x.numel())
        self.seg = None
        self.max_seg = self.n
        if seg is not None:
            seg_h = np.ascontiguousarray(seg, dtype=np.int64)
            self.max_seg = int(np.max(np.diff(seg_h))) if seg_h.size > 1 else 0
            self.seg = torch.from_numpy(seg_h).to(self.device)
        self._scratch = None
        self._logp = None
        self.evals = torch.zeros(1, dtype=torch.int64, device=self.device)
        self.lib = _lib.lib()
        if sort and self.n >= 64:
            # one-time layout step: trials ordered by |rt| (within segments) so that the lanes of a warp
            # take the same series branch; the per-set sums are order-independent up to rounding
            n_seg = 0 if self.seg is None else int(self.seg.numel()) - 1
            nbytes = int(self.lib.hddm_sort_trials_scratch(self.n, n_seg))
            scratch = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
            xs = torch.empty_like(self.x)
            stream = torch.cuda.current_stream(self.device).cuda_stream
            _lib.check(self.lib.hddm_sort_trials_dev(
                self.x.data_ptr(), self.n, self.seg.data_ptr() if self.seg is not None else None, n_seg,
                xs.data_ptr(), scratch.data_ptr(), nbytes, C.c_void_p(stream)))
            torch.cuda.current_stream(self.device).synchronize()
            self.x = xs

    def _ensure(self, P):
        torch = self.torch
        need = int(self.lib.hddm_wiener_like_batch_scratch(self.max_seg, P))
        if self._scratch is None or self._scratch.numel() < need:
            self._scratch = torch.empty(max(need, 1), dtype=torch.float64, device=self.device)
        if self._logp is None or self._logp.numel() != P:
            self._logp = torch.empty(P, dtype=torch.float64, device=self.device)

    def loglik(self, params_dev, out=None, err=1e-4, n_st=2, n_sz=2, use_adaptive=1, simps_err=1e-3):
        """Enqueue one batched launch on torch's current stream.  params_dev: float64 CUDA tensor
        [9, P] (v, sv, a, z, sz, t, st, p_outlier, w_outlier).  Returns the device tensor logp[P]."""
        torch = self.torch
        assert params_dev.is_cuda and params_dev.dtype == torch.float64 and params_dev.is_contiguous()
        P = int(params_dev.shape[1])
        self._ensure(P)
        out = self._logp if out is None else out
        stream = torch.cuda.current_stream(self.device).cuda_stream
        rc = self.lib.hddm_wiener_like_batch_dev(
            self.x.data_ptr(), self.n, self.seg.data_ptr() if self.seg is not None else None, self.max_seg,
            params_dev.data_ptr(), P, err, n_st, n_sz, int(bool(use_adaptive)), simps_err, out.data_ptr(),
            self._scratch.data_ptr(), self.evals.data_ptr(), C.c_void_p(stream))
        _lib.check(rc)
        return out


def gather_loglik(local_logp, n_sets, group=None):
    """All-gather the per-rank blocks (contiguous shard_bounds layout) into the full vector [n_sets].
    Works for CUDA tensors over NCCL and CPU tensors over gloo."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local_logp
    world = dist.get_world_size(group)
    sizes = shard_sizes(n_sets, world)
    width = max(sizes)
    if min(sizes) == width:  # equal shards: one pre-sized all-gather straight into the result, no pad, no cat
        gathered = local_logp.new_empty(n_sets)
        dist.all_gather_into_tensor(gathered, local_logp.contiguous(), group=group)
        return gathered
    padded = local_logp.new_zeros(width)
    padded[:local_logp.numel()] = local_logp
    gathered = local_logp.new_empty(world * width)
    dist.all_gather_into_tensor(gathered, padded, group=group)
    return torch.cat([gathered[r * width:r * width + sizes[r]] for r in range(world)])


def gather_rows(local, n_rows, group=None):
    """All-gather row blocks (contiguous shard_bounds layout over the first axis) of a 2-D tensor into
    the full [n_rows, cols] table: per-draw statistics, quantile probabilities."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size(group) == 1:
        return local
    world = dist.get_world_size(group)
    sizes = shard_sizes(n_rows, world)
    width, cols = max(sizes), int(local.shape[1])
    if min(sizes) == width:
        gathered = local.new_empty((n_rows, cols))
        dist.all_gather_into_tensor(gathered, local.contiguous(), group=group)
        return gathered
    padded = local.new_zeros((width, cols))
    padded[:local.shape[0]] = local
    gathered = local.new_empty((world * width, cols))
    dist.all_gather_into_tensor(gathered, padded, group=group)
    return torch.cat([gathered[r * width:r * width + sizes[r]] for r in range(world)])


def wiener_like_batch_sharded(x, params, evaluate=None, group=None, **numerics):
    """Evaluate P parameter sets (params float64 [9, P], host) against x with the sets sharded over
    the ranks of `group`; every rank returns the full float64[P] vector.

    `evaluate(x, params_block) -> tensor[block]` defaults to the CUDA path (DeviceSweep on this
    rank's GPU); tests pass a stand-in to exercise the partition/gather logic over gloo on CPU."""
    import torch
    import torch.distributed as dist
    params = np.ascontiguousarray(params, dtype=np.float64)
    P = params.shape[1]
    if dist.is_available() and dist.is_initialized():
        world, rank = dist.get_world_size(group), dist.get_rank(group)
    else:
        world, rank = 1, 0
    lo, hi = shard_bounds(P, world, rank)
    block = np.ascontiguousarray(params[:, lo:hi])
    if evaluate is None:
        sweep = DeviceSweep(x)
        nm = dict(HDDM_DEFAULTS, **numerics)
        if hi > lo:
            local = sweep.loglik(torch.from_numpy(block).to(sweep.device), **nm).clone()
        else:
            local = torch.empty(0, dtype=torch.float64, device=sweep.device)
    else:
        local = evaluate(x, block)
    return gather_loglik(local, P, group)
