"""Flat-buffer Adam with the gradient-norm clip folded in.

Replaces, per loss term, `clip_grad_norm_(params, max_norm)` + `torch.optim.Adam.step()`
(reference: iit/model_pairs/iit_behavior_model_pair.py:60-64, base_model_pair.py:176-180,229) by two
kernel launches over one contiguous fp32 buffer: a squared-norm reduction and an Adam update that
applies the clip coefficient on the fly. Parameters and gradients are re-pointed to views of flat
buffers (values preserved), which is also what the data-parallel all-reduce sends in one message.
"""
from __future__ import annotations

import torch

from .. import ops


class FlatAdam(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        params = [p for p in params if p.requires_grad]
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps))
        if not params:
            raise ValueError("FlatAdam: no parameters")
        dev = params[0].device
        if dev.type != "cuda":
            raise RuntimeError("FlatAdam runs on CUDA parameters only (no CPU fallback)")
        self._params = params
        self._offsets = {}
        # 16-byte aligned segments so every view can be read with vector loads
        offs, total = [], 0
        for p in params:
            if p.dtype != torch.float32:
                raise TypeError("FlatAdam: fp32 master parameters expected")
            offs.append(total)
            total += (p.numel() + 3) // 4 * 4
        self.flat_p = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_g = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_m = torch.zeros(total, dtype=torch.float32, device=dev)
        self.flat_v = torch.zeros(total, dtype=torch.float32, device=dev)
        self._sq = torch.zeros(1, dtype=torch.float64, device=dev)
        # step count and learning rate live on the device so that a captured CUDA graph of step() stays valid
        self.step_dev = torch.zeros(1, dtype=torch.int64, device=dev)
        self.lr_dev = torch.full((1,), float(lr), dtype=torch.float32, device=dev)
        self._lr_host = float(lr)
        self.post_step = []   # callables run after every update (e.g. the model's refresh of its bf16 filter copies)
        with torch.no_grad():
            for p, o in zip(params, offs):
                self._offsets[id(p)] = o
                pv = self._view_like(self.flat_p, o, p)
                pv.copy_(p)
                p.data = pv
                p.grad = self._view_like(self.flat_g, o, p)
                p._iit_direct_grad = p.grad     # backward kernels may add into this view in place (ops.direct_grad)

    @staticmethod
    def _view_like(flat, off, p):
        """View of flat[off: off+numel] with p's shape and (dense) strides, e.g. channels-last filters."""
        return torch.as_strided(flat, p.shape, p.stride(), off) if _dense(p) else flat[off: off + p.numel()].view(p.shape)

    def offset_of(self, p) -> int:
        """Element offset of parameter `p` (and of its gradient) in the flat buffers; segments follow parameter order."""
        return self._offsets[id(p)]

    def zero_grad(self, set_to_none: bool = False):
        self.flat_g.zero_()

    def grad_sqnorm(self) -> torch.Tensor:
        self._sq.zero_()
        ops.sqnorm_(self.flat_g, self._sq)
        return self._sq

    @property
    def step_count(self) -> int:
        return int(self.step_dev.item())

    def sync_lr(self):
        """Push a host-side learning-rate change (e.g. from a scheduler) to the device scalar. Call outside
        graph capture; step() calls it when not capturing."""
        lr = float(self.param_groups[0]["lr"])
        if lr != self._lr_host:
            self.lr_dev.fill_(lr)
            self._lr_host = lr

    def state_snapshot(self):
        return [t.clone() for t in (self.flat_p, self.flat_m, self.flat_v, self.step_dev)]

    def state_restore(self, snap):
        for t, s in zip((self.flat_p, self.flat_m, self.flat_v, self.step_dev), snap):
            t.copy_(s)
        for p in self._params:
            torch.autograd.graph.increment_version(p)

    @torch.no_grad()
    def step(self, closure=None, max_norm=None):
        g = self.param_groups[0]
        if not ops.capturing():
            self.sync_lr()
        sq = self.grad_sqnorm() if max_norm else None
        ops.adam_step_dev_(self.flat_p, self.flat_g, self.flat_m, self.flat_v, self.lr_dev, g["betas"][0], g["betas"][1],
                           g["eps"], self.step_dev, sq, max_norm)
        for p in self._params:  # the kernel wrote through raw pointers: tell autograd / filter caches
            torch.autograd.graph.increment_version(p)
        for cb in self.post_step:
            cb()


def _dense(p) -> bool:
    """True when p's strides are a permutation of a dense layout (so as_strided reproduces it)."""
    sizes_strides = sorted(((st, sz) for sz, st in zip(p.shape, p.stride()) if sz > 1), reverse=True)
    expect = 1
    for st, sz in reversed(sizes_strides):
        if st != expect:
            return False
        expect *= sz
    return True
