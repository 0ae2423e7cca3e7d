"""Stop-gradient training: activations of the nodes outside the circuit are scaled down in the forward pass and their
gradients zeroed in the backward pass (API mirror of the reference's iit/model_pairs/stop_grad_pair.py:8-152).

`StopGradHookedModel` wraps the LL model; calling it runs the model with
  * forward hooks `act / scale` on the whole hook point of every node outside the circuit (:37-44, :84-87), and
  * backward hooks that zero the gradient inside each such node's index (:63-77, :88-91),
attached with `reset_hooks_end=False` (:92). As in the reference they therefore STAY attached after the call and also act
on the next `run_with_cache` -- the source run of the following intervention -- whose end then removes them (hooks carry
the context level they were added at, and `run_with_cache` clears its own level). Everything else (`run_with_cache`,
`run_with_hooks`, `parameters`, `train`, ...) is forwarded to the wrapped model untouched (:30-35).
The device work is this library's: `iit_div_scalar` for the scale, `iit_patch_bwd` for the zeroed gradient box."""
from __future__ import annotations

from typing import Any

import torch as t

from iit_b200 import ops
from iit_b200.model_pairs.freeze_model_pair import FreezedModelPair
from iit_b200.model_pairs.nodes import LLNode
from iit_b200.utils import node_picker


class _StopGradSiteFn(t.autograd.Function):
    @staticmethod
    def forward(ctx, act, scale, indices):
        ctx.indices = indices
        if scale:
            return ops.div_scalar(act.detach(), scale)
        return act.detach().clone(memory_format=t.preserve_format)

    @staticmethod
    def backward(ctx, g):
        g = g.clone(memory_format=t.preserve_format)
        for index in ctx.indices:
            ops.zero_box_(g, index)
        return g, None, None


class StopGradHookedModel:
    def __init__(self, model, params_not_in_circuit, nodes_not_in_circuit, post_nodes_not_in_circuit, scale=1e6,
                 use_forward_hooks=True):
        self.model = model
        self.params_not_in_circuit = params_not_in_circuit
        self.nodes_not_in_circuit = nodes_not_in_circuit
        self.post_nodes_not_in_circuit = post_nodes_not_in_circuit
        self.scale = scale
        self.use_forward_hooks = use_forward_hooks

    def __getattr__(self, name: str) -> Any:
        model = self.__dict__.get("model")
        if model is not None and hasattr(model, name):
            return getattr(model, name)
        raise AttributeError(f"'{type(self).__name__}' object has no attribute '{name}'")

    @staticmethod
    def make_ln_hook(ll_node: LLNode, scale):
        """`act / scale` on its own (reference :37-44), differentiable in the ordinary way."""
        def hook_fn(hook_point_out, hook):
            return ops.div_scalar(hook_point_out, scale)

        return hook_fn

    def make_zero_grad_hook(self, ll_node: LLNode):
        """A gradient hook that zeroes `ll_node`'s index (reference :63-77), on a copy of the gradient."""
        def hook_fn(grad, hook):
            return ops.zero_box_(grad.clone(memory_format=t.preserve_format), ll_node.index)

        return hook_fn

    def make_site_hook(self, name: str):
        """Everything the reference attaches to hook point `name`, as ONE forward hook: the scale (when forward hooks are
        on and `name` is a post node) and the zero-gradient boxes of all its nodes outside the circuit. One function
        instead of a forward hook plus backward hooks because of how transformer_lens wires them: its backward hook is a
        full backward hook of the HookPoint MODULE, and what the hook returns (the edited gradient w.r.t. the module's
        output, i.e. of the scaled activation) REPLACES the gradient w.r.t. the module's input -- the 1/scale factor of the
        division is skipped (_StopGradSiteFn.backward). With every node outside the circuit zeroed whole, as for the IOI
        correspondence, both readings give zero; this keeps the as-written behaviour for partial boxes too."""
        scale = self.scale if (self.use_forward_hooks and any(n.name == name for n in self.post_nodes_not_in_circuit)) else None
        indices = tuple(n.index for n in self.nodes_not_in_circuit if n.name == name)

        def hook_fn(hook_point_out, hook):
            return _StopGradSiteFn.apply(hook_point_out, scale, indices)

        return hook_fn

    def forward(self, x, **kwargs):
        self.model.reset_hooks()
        names = []
        for n in (*self.post_nodes_not_in_circuit, *self.nodes_not_in_circuit):
            if n.name not in names:
                names.append(n.name)
        return self.model.run_with_hooks(x, fwd_hooks=[(name, self.make_site_hook(name)) for name in names],
                                         reset_hooks_end=False, **kwargs)

    def __call__(self, *args, **kwargs):
        return self.forward(*args, **kwargs)


class StopGradModelPair(FreezedModelPair):
    def __init__(self, hl_model, ll_model, corr, training_args=None):
        defaults = {
            "batch_size": 256,
            "lr": 0.001,
            "num_workers": 0,
            "use_single_loss": False,
            "iit_weight": 1.0,
            "behavior_weight": 1.0,
            "scale": 1e6,
            "use_ln_hooks": True,
        }
        targs = {**defaults, **(training_args or {})}
        super().__init__(hl_model, ll_model, corr=corr, training_args=targs)
        self.ll_model = StopGradHookedModel(
            ll_model,
            node_picker.get_params_not_in_circuit(corr, ll_model),
            node_picker.get_nodes_not_in_circuit(ll_model, corr),
            node_picker.get_post_nodes_not_in_circuit(ll_model, corr),
            scale=targs["scale"],
            use_forward_hooks=targs["use_ln_hooks"],
        )
        self.wandb_method = "stop grads"
