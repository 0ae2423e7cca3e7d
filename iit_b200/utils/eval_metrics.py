"""Metrics of the ablation evaluations (API mirror of the reference's iit/utils/eval_metrics.py:4-33)."""
from __future__ import annotations

import ctypes as C

import torch

from iit_b200 import ops
from iit_b200._lib import call


def _is_pmf(x: torch.Tensor) -> bool:
    s = x.sum(dim=-1)
    return bool(torch.allclose(s, torch.ones_like(s)))


def kl_div(a: torch.Tensor, b: torch.Tensor, label_idx) -> torch.Tensor:
    """Per-sample KL(b || a) at the label positions; rows that are not probability vectors go through a softmax first
    (eval_metrics.py:4-20). One kernel (iit_kl_div_rows) does the softmaxes and the sum."""
    a_sel, b_sel = a[label_idx.as_index], b[label_idx.as_index]
    ops._require_cuda(a_sel, b_sel)
    if a_sel.shape != b_sel.shape:
        raise ValueError(f"kl_div: shapes differ: {tuple(a_sel.shape)} vs {tuple(b_sel.shape)}")
    lead = a_sel.shape[:-1]
    a2 = a_sel.reshape(-1, a_sel.shape[-1]).float()
    b2 = b_sel.reshape(-1, b_sel.shape[-1]).float()
    if a2.stride(1) != 1:
        a2 = a2.contiguous()
    if b2.stride(1) != 1:
        b2 = b2.contiguous()
    out = torch.empty(a2.shape[0], dtype=torch.float32, device=a2.device)
    call("iit_kl_div_rows", ops._p(a2), ops._p(b2), a2.shape[0], a2.shape[1], a2.stride(0), b2.stride(0),
         int(_is_pmf(a2)), int(_is_pmf(b2)), ops._p(out), ops._stream())
    return out.view(lead)


def accuracy_affected(a: torch.Tensor, b: torch.Tensor, label_unchanged: torch.Tensor, label_idx) -> torch.Tensor:
    """Fraction of the samples whose label changed for which argmax(a) != argmax(b) (eval_metrics.py:22-33)."""
    a_sel, b_sel = a[label_idx.as_index], b[label_idx.as_index]
    a_lab = ops.argmax_rows(a_sel.reshape(-1, a_sel.shape[-1]).float())
    b_lab = ops.argmax_rows(b_sel.reshape(-1, b_sel.shape[-1]).float())
    out_changed = (a_lab != b_lab).view(label_unchanged.shape)
    changed_result = out_changed.float() * (~label_unchanged).float()
    return (changed_result.sum() / (~label_unchanged).sum()).cpu()
