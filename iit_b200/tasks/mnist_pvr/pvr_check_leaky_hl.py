"""Leaky PVR HL model used by the IIA sweep: 12 nodes `hook_{i}_leaked_to_{j}` that interchange
variable i at the HL while quadrant j is patched at the LL
(API mirror of the reference's iit/tasks/mnist_pvr/pvr_check_leaky_hl.py:10-109)."""
from __future__ import annotations

import torch as t

from iit_b200 import ops
from iit_b200.hooks import HookedRootModule, HookPoint
from iit_b200.model_pairs.nodes import HLNode, LLNode
from iit_b200.tasks.mnist_pvr.pvr_hl import MNIST_CLASS_MAP, _HOOKS, _probe_shape, quadrant_indices, var_index

_Q = ["tl", "tr", "bl", "br"]


class MNIST_PVR_Leaky_HL(HookedRootModule):
    def __init__(self, class_map=MNIST_CLASS_MAP, device=None):
        super().__init__()
        self.hook_tl = HookPoint()
        self.hook_tr = HookPoint()
        self.hook_bl = HookPoint()
        self.hook_br = HookPoint()
        self.leaky_hooks = {}
        for i in _Q:
            for j in _Q:
                if i != j:
                    node = HLNode(f"hook_{i}_leaked_to_{j}", 10, None)
                    hp = HookPoint()
                    self.leaky_hooks[node] = hp
                    setattr(self, node.name, hp)
        self.setup()

    def is_categorical(self):
        return True

    def get_idx_to_intermediate(self, name):
        i = var_index(name)
        return lambda intermediate_vars: intermediate_vars[:, i]

    def forward(self, args):
        _, _, intermediate_data = args
        if not any(h.has_hooks() for h in self.hook_dict.values()):
            return ops.hl_pvr(intermediate_data, None, -1)
        v = [getattr(self, h)(intermediate_data[:, i]) for i, h in enumerate(_HOOKS)]
        for node, hp in self.leaky_hooks.items():
            i = var_index(node.name)
            v[i] = hp(v[i])
        return ops.hl_pvr(t.stack(v, dim=1), None, -1)

    def interchange(self, base_input, ablation_input, hl_node) -> t.Tensor:
        return ops.hl_pvr(base_input[2], ablation_input[2], var_index(hl_node.name))

    def interchange_ce(self, base_input, ablation_input, hl_node, ll_output):
        return ops.hl_cross_entropy(ll_output, ("pvr", base_input[2], ablation_input[2], var_index(hl_node.name)))


def get_corr(mode, hook_point, model, input_shape):
    if mode != "q":
        raise NotImplementedError(mode)
    shape = _probe_shape(model, hook_point, input_shape)
    idxs = quadrant_indices(shape[2] // 2)
    corr = {}
    for i in _Q:
        for j in _Q:
            if i != j:
                corr[HLNode(f"hook_{i}_leaked_to_{j}", 10, None)] = {LLNode(name=hook_point, index=idxs[_Q.index(j)])}
    return corr
