"""Data parallelism over (base, source) pairs: one process per GPU, NCCL over NVLink.

The reference has no distributed code; pairs are independent units, so the
only exchange is the gradient all-reduce before the clip + Adam step of every loss term, plus an
integer all-reduce of IIA counts in evaluation. Node sampling needs no communication: every rank
owns the same `np.random.default_rng(seed)` stream and draws in the same order, so all ranks patch
the same node within one global batch (as the reference does within one batch).
BatchNorm statistics stay per-rank (each rank's shard is its own "batch")."""
from __future__ import annotations

import os

import torch
import torch.distributed as dist


def init_from_env(backend=None):
    """torchrun-style rendezvous (RANK / LOCAL_RANK / WORLD_SIZE / MASTER_*). Returns (rank, world, local_rank)."""
    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if world > 1 and not dist.is_initialized():
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("MASTER_PORT", "29500")
        if backend is None:
            backend = "nccl" if torch.cuda.is_available() else "gloo"
        if backend == "nccl":
            torch.cuda.set_device(local)
            dist.init_process_group(backend, rank=rank, world_size=world, device_id=torch.device("cuda", local))
        else:
            dist.init_process_group(backend, rank=rank, world_size=world)
    return rank, world, local


class DataParallel:
    def __init__(self, group=None):
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rank = dist.get_rank(group) if dist.is_initialized() else 0

    def shard(self, inputs):
        """This rank's equal slice of a global batch (tuple of batched tensors); ragged tails are dropped
        identically on every rank so that mean-of-shard-means equals the global mean."""
        n = inputs[0].shape[0] // self.world
        return tuple(t[self.rank * n:(self.rank + 1) * n] for t in inputs)

    def allreduce_grads(self, optimizer, ll_model):
        """Average gradients across ranks, before clip_grad_norm_ so the clip sees the global norm
        (reference order: backward -> clip -> step, iit_behavior_model_pair.py:60-64)."""
        if self.world == 1:
            return
        flat = getattr(optimizer, "flat_g", None)
        if flat is not None:  # FlatAdam: one message
            dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=self.group)
            flat.mul_(1.0 / self.world)
            return
        for p in ll_model.parameters():
            if p.grad is not None:
                dist.all_reduce(p.grad, op=dist.ReduceOp.SUM, group=self.group)
                p.grad.mul_(1.0 / self.world)

    def allreduce_counts(self, counts: torch.Tensor) -> torch.Tensor:
        """Sum int64 (hits, denominator) counts over ranks: reduce counts, then divide."""
        if self.world > 1:
            dist.all_reduce(counts, op=dist.ReduceOp.SUM, group=self.group)
        return counts

    def allgather_counts(self, counts: torch.Tensor) -> torch.Tensor:
        """[entries, batches, 2] per rank -> [entries, world * batches, 2]: the sweep keeps per-batch counts because the
        reference averages per-batch ratios (eval_causality.py:85-88)."""
        if self.world == 1:
            return counts
        parts = [torch.empty_like(counts) for _ in range(self.world)]
        dist.all_gather(parts, counts.contiguous(), group=self.group)
        return torch.cat(parts, dim=1)

    def max_over_ranks(self, seconds: float, device) -> float:
        t = torch.tensor([seconds], dtype=torch.float64, device=device)
        if self.world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX, group=self.group)
        return float(t.item())

    def barrier(self):
        if self.world > 1:
            dist.barrier(group=self.group)
