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

    # ---- gradient averaging ---------------------------------------------------------------------------------------
    # The flat gradient buffer is ordered like the parameters (stem first, head last) while the backward fills it from
    # the END. A model may name a "tap": a hook point whose gradient is computed only after every parameter from
    # `first_tail_param` on has received its gradient (ResNet-18: the output of layer3, behind which sit layer4 and the
    # head = 75 % of the parameters). When the tap's gradient arrives, the tail of the buffer is all-reduced at once on
    # NCCL's stream, under the rest of the backward; only the (small) head part is reduced after it. Two messages, the
    # larger one hidden. NCCL averages in the collective (ReduceOp.AVG): no separate scaling kernel.
    def _average(self, t, async_op=False):
        if dist.get_backend(self.group) == "nccl":
            return dist.all_reduce(t, op=dist.ReduceOp.AVG, group=self.group, async_op=async_op)
        w = dist.all_reduce(t, op=dist.ReduceOp.SUM, group=self.group, async_op=False)   # gloo has no AVG
        t.mul_(1.0 / self.world)
        return None if async_op else w

    def _attach(self, optimizer, ll_model):
        self._attached = (optimizer, ll_model)
        self._tail = None          # (offset of the tail bucket, pending Work or None)
        tap = getattr(ll_model, "dp_bucket_tap", None)
        offset_of = getattr(optimizer, "offset_of", None)
        if tap is None or offset_of is None or getattr(optimizer, "flat_g", None) is None:
            return
        hook_name, first_tail_param = tap
        off = offset_of(dict(ll_model.named_parameters())[first_tail_param])
        flat = optimizer.flat_g

        def tail_ready(grad, hook):
            if self._tail is None:      # once per backward: every gradient from `off` on is final
                self._tail = (off, self._average(flat[off:], async_op=True))
            return None

        ll_model.hook_dict[hook_name].add_hook(tail_ready, dir="bwd", is_permanent=True)

    def allreduce_grads(self, optimizer, ll_model):
        """Average gradients across ranks, before clip_grad_norm_ so the clip sees the global norm
        (reference order: backward -> clip -> step, iit_behavior_model_pair.py:60-64)."""
        if self.world == 1:
            return
        if getattr(self, "_attached", None) is None or self._attached[0] is not optimizer:
            self._attach(optimizer, ll_model)
        flat = getattr(optimizer, "flat_g", None)
        if flat is not None:  # FlatAdam: one message, or two when the tail bucket left early
            tail, self._tail = getattr(self, "_tail", None), None
            if tail is None:
                self._average(flat)
                return
            off, work = tail
            self._average(flat[:off])
            if work is not None:
                work.wait()       # the current stream waits for the tail's collective
            return
        for p in ll_model.parameters():
            if p.grad is not None:
                self._average(p.grad)

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
