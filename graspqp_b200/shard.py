"""Object sharding across the GPUs of one box.

The reference has no distributed code: scripts/fit_all.sh starts one process per group of assets, each optimising its
own objects.  The same partition here is one process per GPU (torchrun), objects dealt round-robin, and NO data-path
collective by default.  Two quantities of an iteration span the whole batch of a process in the reference — the RMS
normaliser (g**2).mean(0) of MalaStar.try_step (core/optimizer.py:231) and qpth's batch-wide exit test — so a sharded
run only reproduces a single-process run over all objects bit for bit if those two are exchanged: `GlobalStats` does
that with two tiny all-reduces per iteration (D+1 floats, 128 int32).  At export time `gather_results` all-gathers
(pose, energy, contact indices).  Works with NCCL (device tensors) and with gloo (host buffers: the CPU tests).
"""
import numpy as np


def objects_of_rank(n_objects, rank, world):
    """Round-robin: rank r optimises objects r, r+world, ...  -> list of global object indices."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world {world}")
    return list(range(rank, n_objects, world))


def owner_of_object(obj, world):
    return obj % world


def _as_tensor(a):
    import torch
    return a if isinstance(a, torch.Tensor) else torch.from_numpy(a)  # shares memory with the numpy buffer


class GlobalStats:
    """The two batch-wide statistics of an iteration, reduced over all ranks (engine.FitEngine(collectives=...))."""

    def __init__(self, group=None):
        import torch.distributed as dist
        self.dist, self.group = dist, group
        self.calls = 0

    def reduce_rms(self, grad2_mean, n_rows):
        """grad2_mean (D): mean over this rank's rows of grad**2  ->  mean over the rows of all ranks, in place."""
        import torch
        t = _as_tensor(grad2_mean)
        buf = torch.cat([t.detach().to(torch.float64) * float(n_rows),
                         torch.tensor([float(n_rows)], dtype=torch.float64, device=t.device)])
        self.dist.all_reduce(buf, op=self.dist.ReduceOp.SUM, group=self.group)
        t.copy_((buf[:-1] / buf[-1]).to(t.dtype))
        self.calls += 1

    def reduce_qp_header(self, header_i32):
        """header (128 int32): per interior-point iteration the float bits of the batch maximum of the best residuals
        (non-negative floats order like their bit patterns) and the any-improved flags -> element-wise max, in place."""
        t = _as_tensor(header_i32)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX, group=self.group)
        self.calls += 1


def gather_results(pose, energy, idx, n_objects, batch_each, rank, world, group=None):
    """All-gather the final grasps of every rank and return them in GLOBAL object order:
    (pose (n_objects*B, D), energy (n_objects*B), idx (n_objects*B, n_c)) as torch tensors on the inputs' device.
    Ranks may own different numbers of objects; shorter ranks are padded for the collective."""
    import torch
    import torch.distributed as dist
    pose, energy, idx = _as_tensor(pose), _as_tensor(energy), _as_tensor(idx)
    per_rank = [len(objects_of_rank(n_objects, r, world)) for r in range(world)]
    rows_max = max(per_rank) * batch_each
    D, n_c = pose.shape[1], idx.shape[1]
    packed = torch.zeros(rows_max, D + 1 + n_c, dtype=torch.float32, device=pose.device)
    n = pose.shape[0]
    assert n == per_rank[rank] * batch_each, "rows do not match this rank's objects"
    packed[:n, :D] = pose
    packed[:n, D] = energy
    packed[:n, D + 1:] = idx.to(torch.float32)  # contact indices < 2**24: exact in fp32
    out = [torch.empty_like(packed) for _ in range(world)]
    dist.all_gather(out, packed, group=group)
    full = torch.zeros(n_objects * batch_each, D + 1 + n_c, dtype=torch.float32, device=pose.device)
    for r in range(world):
        for k, o in enumerate(objects_of_rank(n_objects, r, world)):
            full[o * batch_each:(o + 1) * batch_each] = out[r][k * batch_each:(k + 1) * batch_each]
    return full[:, :D], full[:, D], full[:, D + 1:].to(torch.int32)
