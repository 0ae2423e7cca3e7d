"""Counterfactual inputs for the causality sweep, generated on the device.

Replaces `MNIST_PVR_Dataset.patch_batch_at_hl` / `patch_at_hl_idx` / `get_idx_and_intermediate`
(reference: iit/tasks/mnist_pvr/dataset.py:125-195), which, per sample, re-draws random digits on the host until one has
a different class, converts it from PIL and pastes it into the quadrant. Here the digits are a device-resident bank
sorted by label; one kernel draws, for every sample, a bank item of another class (ops.pick_other_class), one kernel
pastes the digits (ops.gather_box_), and the new label is the HL model evaluated on the new intermediate variables.
Same distribution (uniform over the digits whose class differs); the random stream is this library's, not numpy's."""
from __future__ import annotations

import torch as t

from iit_b200 import ops
from iit_b200.tasks.mnist_pvr.pvr_hl import var_index


class DigitBank:
    def __init__(self, images: t.Tensor, labels: t.Tensor, n_classes: int = 10, seed: int = 0):
        """images [M, C, h, w] (C = 3, or 1: repeated to 3 like dataset.py:144-145), labels [M] int64."""
        if images.dim() != 4 or labels.dim() != 1 or images.shape[0] != labels.shape[0]:
            raise ValueError("DigitBank: images [M,C,h,w] and labels [M] expected")
        if not images.is_cuda:
            raise RuntimeError("DigitBank lives on the GPU (no CPU fallback exists)")
        if images.shape[1] == 1:
            images = images.repeat(1, 3, 1, 1)
        order = t.argsort(labels.to(images.device), stable=True)
        self.images = images[order].contiguous()
        self.labels = labels.to(images.device)[order].contiguous()
        counts = t.bincount(self.labels, minlength=n_classes)
        self.class_start = t.cat([t.zeros(1, dtype=t.int64, device=images.device), t.cumsum(counts, 0)]).contiguous()
        self.n_classes = n_classes
        self.seed = seed
        self._draws = 0

    def get_idx_and_intermediate(self, hl_node, input_shape):
        """(row offset, column offset, quadrant height, quadrant width, index of the variable): dataset.py:176-195."""
        h, w = input_shape[2], input_shape[3]
        i = var_index(getattr(hl_node, "name", str(hl_node)))
        return (i // 2) * (h // 2), (i % 2) * (w // 2), h // 2, w // 2, i

    def patch_batch_at_hl(self, batch, intermediate_vars, hl_node, labels=None):
        """(new_x, new_labels, new_intermediate_vars) as stacked device tensors; `batch` / `intermediate_vars` may be
        the stacked tensors or the per-sample lists the reference passes."""
        x = t.stack(list(batch)) if not isinstance(batch, t.Tensor) else batch
        v = t.stack(list(intermediate_vars)) if not isinstance(intermediate_vars, t.Tensor) else intermediate_vars
        r0, c0, qh, qw, i = self.get_idx_and_intermediate(hl_node, x.shape)
        if (qh, qw) != tuple(self.images.shape[2:]):
            raise ValueError(f"quadrant {qh}x{qw} does not match the bank's digits {tuple(self.images.shape[2:])}")
        self._draws += 1
        idx, lab = ops.pick_other_class(v[:, i].contiguous(), self.class_start, self.images.shape[0],
                                        self.seed * 1_000_003 + self._draws)
        new_x = x.clone()
        ops.gather_box_(new_x, self.images.to(new_x.dtype), idx, r0, c0)
        new_v = v.clone()
        new_v[:, i] = lab
        new_y = ops.hl_pvr(new_v, None, -1)          # make_label_from_intermediate (dataset.py:109-111)
        return new_x, new_y, new_v
