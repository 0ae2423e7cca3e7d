"""High-level IOI model (API mirror of the reference's iit/tasks/ioi/ioi_hl.py:73-130).

The reference materialises [B,S,V] logits (zeros + two index_puts + cumsum; 823 MB at batch 256) although
every consumer reads only argmax(out[:, -1]). Here:
  * `forward(args)` returns the same dense tensor from one kernel (API compatibility);
  * `interchange(base, ablation, hl_node)` returns the dense output of the base run with one hook's value taken
    from the ablation run (what do_intervention's HL half computes, base_model_pair.py:80,89-91);
  * `interchange_last_argmax(...)` returns argmax(out[:, -1, :]) directly from the sparse evaluation.
Hooks on the internal hook points are not supported (their values exist only inside the kernel)."""
from __future__ import annotations

import torch as t

from iit_b200 import ops
from iit_b200.hooks import HookedRootModule, HookPoint
from iit_b200.tasks.hl_model import HLModel

_NODES = ("all_nodes_hook", "hook_duplicate", "hook_s_inhibition", "hook_name_mover")


class IOI_HL(HookedRootModule, HLModel):
    def __init__(self, d_vocab, names):
        super().__init__()
        self.all_nodes_hook = HookPoint()
        self.hook_duplicate = HookPoint()
        self.hook_previous = HookPoint()
        self.hook_s_inhibition = HookPoint()
        self.hook_name_mover = HookPoint()
        self.d_vocab = int(d_vocab)
        self.register_buffer("names", t.sort(t.as_tensor(names).to(t.int64).flatten()).values)  # sorted for the kernel
        self.setup()

    def is_categorical(self):
        return True

    def _no_hooks(self):
        if any(h.has_hooks() for h in self.hook_dict.values()):
            raise NotImplementedError("IOI_HL evaluates inside one kernel: use interchange() / interchange_last_argmax() "
                                      "to swap a node's value instead of hooking its hook point")

    def _tokens(self, inp):
        x = inp[0]
        return x[None, ...] if x.dim() == 1 else x

    def forward(self, args, verbose=False):
        self._no_hooks()
        x = args[0]
        out = ops.hl_ioi_dense(self._tokens(args), None, None, self.names, self.d_vocab)
        return out[0] if x.dim() == 1 else out

    def _node_name(self, hl_node):
        name = str(hl_node)
        if name not in _NODES:
            raise KeyError(f"unknown IOI HL node {name}")
        return name

    def interchange(self, base_input, ablation_input, hl_node) -> t.Tensor:
        self._no_hooks()
        return ops.hl_ioi_dense(self._tokens(base_input), self._tokens(ablation_input), self._node_name(hl_node),
                                self.names, self.d_vocab)

    def interchange_last_ce(self, base_input, ablation_input, hl_node, last_logits, classes):
        """(loss, hl_argmax): the intervened HL run's last-position argmax and the mean cross-entropy of `last_logits[:, :classes]`
        against it, in one launch (ops.hl_cross_entropy)."""
        self._no_hooks()
        return ops.hl_cross_entropy(last_logits, ("ioi", self._tokens(base_input), self._tokens(ablation_input),
                                                  self._node_name(hl_node), self.names, self.d_vocab), classes=classes)

    def interchange_last_argmax(self, base_input, ablation_input, hl_node) -> t.Tensor:
        self._no_hooks()
        return ops.hl_ioi_last_argmax(self._tokens(base_input), self._tokens(ablation_input), self._node_name(hl_node),
                                      self.names, self.d_vocab)
