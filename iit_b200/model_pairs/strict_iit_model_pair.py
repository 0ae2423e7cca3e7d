"""IIT + behaviour + strict loss on a sampled LL node outside the circuit
(API mirror of the reference's iit/model_pairs/strict_iit_model_pair.py:5-91).

For the ResNet LL model the reference has no definition of "nodes not in the circuit" (its
constructor needs `ll_model.cfg.n_heads`); here they are the `conv` hook points taken whole that
the correspondence does not touch (iit_b200/utils/node_picker.py), overridable through
training_args["nodes_not_in_circuit"]."""
from __future__ import annotations

import torch as t

from iit_b200.model_pairs.iit_behavior_model_pair import IITBehaviorModelPair
from iit_b200.model_pairs.nodes import LLNode
from iit_b200.utils import node_picker
from iit_b200.utils.metric import MetricStore, MetricStoreCollection, MetricType


class StrictIITModelPair(IITBehaviorModelPair):
    def __init__(self, hl_model, ll_model, corr, training_args=None):
        defaults = {
            "batch_size": 256,
            "lr": 0.001,
            "num_workers": 0,
            "use_single_loss": False,
            "iit_weight": 1.0,
            "behavior_weight": 1.0,
            "strict_weight": 1.0,
            "clip_grad_norm": 1.0,
        }
        super().__init__(hl_model, ll_model, corr=corr, training_args={**defaults, **(training_args or {})})
        override = self.training_args.get("nodes_not_in_circuit")
        self.nodes_not_in_circuit = list(override) if override is not None else \
            node_picker.get_nodes_not_in_circuit(self.ll_model, self.corr)

    @staticmethod
    def make_train_metrics():
        return MetricStoreCollection([MetricStore("train/iit_loss", MetricType.LOSS),
                                      MetricStore("train/behavior_loss", MetricType.LOSS),
                                      MetricStore("train/strict_loss", MetricType.LOSS)])

    def sample_ll_node(self) -> LLNode:
        return self.rng.choice(self.nodes_not_in_circuit)

    def get_strict_loss_over_batch(self, base_input, ablation_input, ll_node, loss_fn):
        """Source run cached at `ll_node`, base run patched there, loss against the BASE label
        (reference :57-74): a node outside the circuit must not change the output."""
        base_x, base_y, _ = base_input
        self.ll_source_run(ablation_input[0], [ll_node])
        out = self.ll_model.run_with_hooks(base_x, fwd_hooks=[(ll_node.name, self.make_ll_ablation_hook(ll_node))])
        if out.dim() > 1 and out.shape[0] > 1:
            out = out.squeeze()
        return loss_fn(out, base_y)

    def _graphed_strict(self, sb, ss, ll_node, loss_fn, optimizer):
        return self._graphed_term(("strict", ll_node.name, repr(ll_node.index)), optimizer, lambda: (
            self.get_strict_loss_over_batch(sb, ss, ll_node, loss_fn) * self.training_args["strict_weight"]))

    def prepare_graphs(self, base_input, ablation_input, loss_fn, optimizer):
        """Capture every (term, node) variant up front so that no capture happens inside a timed / training
        loop. Model and optimizer state are left exactly as found."""
        sb, ss = self._static_inputs(base_input, ablation_input)
        for hl_node in self.corr.keys():
            self._graphed_capture_only(("iit", str(hl_node)), optimizer, lambda n=hl_node: (
                self.get_IIT_loss_over_batch(sb, ss, n, loss_fn) * self.training_args["iit_weight"]))
        for ll_node in self.nodes_not_in_circuit:
            self._graphed_capture_only(("strict", ll_node.name, repr(ll_node.index)), optimizer, lambda n=ll_node: (
                self.get_strict_loss_over_batch(sb, ss, n, loss_fn) * self.training_args["strict_weight"]))
        self._graphed_capture_only(("behavior",), optimizer, lambda: (
            self.get_behaviour_loss_over_batch(sb, loss_fn) * self.training_args["behavior_weight"]))

    def _graphed_capture_only(self, key, optimizer, build_loss):
        graphs = self.__dict__.setdefault("_graphs", {})
        if key not in graphs:
            graphs[key] = self._capture_term(optimizer, build_loss)

    def run_train_step(self, base_input, ablation_input, loss_fn, optimizer):
        if self._use_graphs(optimizer):
            sb, ss = self._static_inputs(base_input, ablation_input)
            optimizer.sync_lr()
            hl_node = self.sample_hl_name()
            iit_loss = self._graphed_iit(sb, ss, hl_node, loss_fn, optimizer)
            ll_node = self.sample_ll_node()
            ll_loss = self._graphed_strict(sb, ss, ll_node, loss_fn, optimizer)
            behavior_loss = self._graphed_behavior(sb, loss_fn, optimizer)
            vals = t.stack([iit_loss, behavior_loss, ll_loss]).tolist()
            return {"train/iit_loss": vals[0], "train/behavior_loss": vals[1], "train/strict_loss": vals[2]}
        single = self.training_args["use_single_loss"]
        hl_node = self.sample_hl_name()
        iit_loss = (self.get_IIT_loss_over_batch(base_input, ablation_input, hl_node, loss_fn)
                    * self.training_args["iit_weight"])
        if not single:
            self.step_on_loss(iit_loss, optimizer)
        ll_node = self.sample_ll_node()  # drawn after the HL node on the same stream (reference :62)
        ll_loss = (self.get_strict_loss_over_batch(base_input, ablation_input, ll_node, loss_fn)
                   * self.training_args["strict_weight"])
        if not single:
            self.step_on_loss(ll_loss, optimizer)
        behavior_loss = self.get_behaviour_loss_over_batch(base_input, loss_fn) * self.training_args["behavior_weight"]
        if not single:
            self.step_on_loss(behavior_loss, optimizer)
        else:
            self.step_on_loss(iit_loss + behavior_loss + ll_loss, optimizer)
        vals = t.stack([iit_loss.detach(), behavior_loss.detach(), ll_loss.detach()]).tolist()
        return {"train/iit_loss": vals[0], "train/behavior_loss": vals[1], "train/strict_loss": vals[2]}
