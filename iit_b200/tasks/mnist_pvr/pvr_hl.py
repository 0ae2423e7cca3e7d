"""High-level MNIST-PVR model and its LL correspondences
(API mirror of the reference's iit/tasks/mnist_pvr/pvr_hl.py:10-134).

`forward` and `interchange` evaluate `(tr, bl, br)[class_map[tl] - 1]` in one kernel (iit_hl_pvr);
hooks on hook_tl/tr/bl/br still work (they see and may replace the [B] int64 variable)."""
from __future__ import annotations

import torch as t

from iit_b200 import ops
from iit_b200.hooks import HookedRootModule, HookPoint
from iit_b200.model_pairs.nodes import HLNode, LLNode
from iit_b200.tasks.hl_model import HLModel
from iit_b200.utils.index import Ix

MNIST_CLASS_MAP = {k: [1, 1, 1, 1, 2, 2, 2, 3, 3, 3][k] for k in range(10)}
_HOOKS = ["hook_tl", "hook_tr", "hook_bl", "hook_br"]


def var_index(name: str) -> int:
    for i, h in enumerate(_HOOKS):
        if h in name:
            return i
    raise NotImplementedError(name)


class MNIST_PVR_HL(HookedRootModule, HLModel):
    def __init__(self, class_map=MNIST_CLASS_MAP, device=None):
        super().__init__()
        if [class_map[i] for i in range(10)] != [1, 1, 1, 1, 2, 2, 2, 3, 3, 3]:
            raise NotImplementedError("the PVR kernel implements the reference's MNIST_CLASS_MAP")
        self.hook_tl = HookPoint()
        self.hook_tr = HookPoint()
        self.hook_bl = HookPoint()
        self.hook_br = HookPoint()
        self.setup()

    def is_categorical(self):
        return True

    def get_idx_to_intermediate(self, name):
        i = var_index(name)
        return lambda intermediate_vars: intermediate_vars[:, i]

    def _hooked_vars(self, intermediate_data):
        hps = [self.hook_tl, self.hook_tr, self.hook_bl, self.hook_br]
        if not any(h.has_hooks() for h in hps):
            return intermediate_data
        return t.stack([h(intermediate_data[:, i]) for i, h in enumerate(hps)], dim=1)

    def forward(self, args):
        _, _, intermediate_data = args
        return ops.hl_pvr(self._hooked_vars(intermediate_data), None, -1)

    def interchange(self, base_input, ablation_input, hl_node) -> t.Tensor:
        """HL output of the base run with `hl_node`'s variable taken from the ablation run."""
        if hl_node.index is not None and not hl_node.index.is_whole():
            raise NotImplementedError("PVR HL nodes are interchanged whole (pvr_hl.py:60-65)")
        return ops.hl_pvr(base_input[2], ablation_input[2], var_index(hl_node.name))


    def interchange_ce(self, base_input, ablation_input, hl_node, ll_output):
        """(loss, hl_output): the intervened HL output and the mean cross-entropy of `ll_output` against it in one launch."""
        if hl_node.index is not None and not hl_node.index.is_whole():
            raise NotImplementedError("PVR HL nodes are interchanged whole (pvr_hl.py:60-65)")
        return ops.hl_cross_entropy(ll_output, ("pvr", base_input[2], ablation_input[2], var_index(hl_node.name)))


hl_nodes = {h: HLNode(h, 10, None) for h in _HOOKS}


def _probe_shape(model, hook_point, input_shape):
    """Like the reference (pvr_hl.py:69-71) this is a real forward of a zeros batch in the model's
    CURRENT mode, so in train mode it moves the BatchNorm running statistics exactly as the
    reference's probe does."""
    dev = next(model.parameters()).device
    with t.no_grad():
        _, cache = model.run_with_cache(t.zeros(input_shape, device=dev), names_filter=[hook_point])
    shape = cache[hook_point].shape
    if shape[2] != shape[3]:
        raise AssertionError(f"Input shape is not square, got {shape}")
    return shape


def quadrant_indices(q):
    return [Ix[None, None, :q, :q], Ix[None, None, :q, q:q * 2], Ix[None, None, q:q * 2, :q],
            Ix[None, None, q:q * 2, q:q * 2]]


def get_corr(mode, hook_point, model, input_shape):
    shape = _probe_shape(model, hook_point, input_shape)
    if mode == "c":
        cs = shape[1] // 4
        idxs = [Ix[None, :cs, None, None]] + [Ix[None, cs * i:cs * (i + 1), None, None] for i in range(1, 4)]
    elif mode == "q":
        idxs = quadrant_indices(shape[2] // 2)
    else:
        raise NotImplementedError(mode)
    return {hl_nodes[h]: {LLNode(hook_point, idxs[i])} for i, h in enumerate(_HOOKS)}
