"""IIT + behaviour training that "freezes" the parameters outside the circuit
(API mirror of the reference's iit/model_pairs/freeze_model_pair.py:8-38).

What the reference does AS WRITTEN, which is what this class reproduces:
  * `params_not_in_circuit` = node_picker.get_params_not_in_circuit(corr, ll_model): parameter names are compared with
    hook-point names, which never match, so it lists every non-embedding parameter slice (reference
    iit/utils/node_picker.py:157-170);
  * `zero_grad_for_not_in_circuit` assigns `.grad` on the temporary `param[idx]` (:28), which does not touch
    `param.grad`: the gradient that reaches the optimizer is unchanged. The only observable effect is the assertion that
    every listed parameter HAS a gradient after the backward (:30), kept here.
So the update equals IITBehaviorModelPair's; the loss is IITBehaviorModelPair's too -- for a transformer LL model that is
CrossEntropyLoss over dim 1 of the [B, S, V] logits (torch's class dimension), not the IOI pair's last-position loss."""
from __future__ import annotations

from iit_b200.model_pairs.iit_behavior_model_pair import IITBehaviorModelPair
from iit_b200.nn.optim import FlatAdam
from iit_b200.utils import node_picker


class FreezedModelPair(IITBehaviorModelPair):
    def __init__(self, hl_model, ll_model, corr, training_args=None):
        defaults = {
            "batch_size": 256,
            "lr": 0.001,
            "num_workers": 0,
            "use_single_loss": False,
            "iit_weight": 1.0,
            "behavior_weight": 1.0,
        }
        super().__init__(hl_model, ll_model, corr=corr, training_args={**defaults, **(training_args or {})})
        self.params_not_in_circuit = node_picker.get_params_not_in_circuit(corr, ll_model)
        self.wandb_method = "freeze_unwanted"

    def zero_grad_for_not_in_circuit(self):
        """No gradient changes (see the module docstring); every listed parameter must have received one."""
        named = dict(self.ll_model.named_parameters())
        for ll_node in self.params_not_in_circuit:
            param = named.get(ll_node.name)
            if param is not None and param.grad is None:
                raise AssertionError(f"{ll_node.name} has no gradient")

    def _use_graphs(self, optimizer) -> bool:
        return False   # the reference's step order (zero_grad, backward, check, clip, step) stays visible to hooks

    def step_on_loss(self, loss, optimizer):
        optimizer.zero_grad()
        loss.backward()
        self.zero_grad_for_not_in_circuit()
        if self.dp is not None:
            self.dp.allreduce_grads(optimizer, self.ll_model)
        if isinstance(optimizer, FlatAdam):
            optimizer.step(max_norm=self.training_args["clip_grad_norm"])
        else:
            self.clip_grad_fn()
            optimizer.step()
