"""IIT + behaviour loss, with the IIA metric: the pair `train.py` uses
(API mirror of the reference's iit/model_pairs/iit_behavior_model_pair.py:6-141)."""
from __future__ import annotations

import torch as t

from iit_b200 import ops
from iit_b200.model_pairs.iit_model_pair import IITModelPair, _f32_ratio
from iit_b200.nn.optim import FlatAdam
from iit_b200.utils.metric import MetricStore, MetricStoreCollection, MetricType


class IITBehaviorModelPair(IITModelPair):
    def __init__(self, hl_model, ll_model, corr, training_args=None):
        defaults = {
            "lr": 0.001,
            "atol": 5e-2,
            "early_stop": True,
            "use_single_loss": False,
            "iit_weight": 1.0,
            "behavior_weight": 1.0,
        }
        super().__init__(hl_model, ll_model, corr=corr, training_args={**defaults, **(training_args or {})})
        self.wandb_method = "iit_and_behavior"

    @property
    def loss_fn(self):
        try:
            categorical = self.hl_model.is_categorical()
        except AttributeError:
            print("WARNING: using default categorical loss function.")
            categorical = True
        # regression HL models: torch's MSELoss on CUDA tensors (no reference task here has one; not a kernel of this library)
        return ops.cross_entropy if categorical else t.nn.MSELoss()

    @staticmethod
    def make_train_metrics():
        return MetricStoreCollection([MetricStore("train/iit_loss", MetricType.LOSS),
                                      MetricStore("train/behavior_loss", MetricType.LOSS)])

    @staticmethod
    def make_test_metrics():
        return MetricStoreCollection([MetricStore("val/iit_loss", MetricType.LOSS),
                                      MetricStore("val/IIA", MetricType.ACCURACY),
                                      MetricStore("val/accuracy", MetricType.ACCURACY)])

    def get_behaviour_loss_over_batch(self, base_input, loss_fn):
        base_x, base_y, _ = base_input
        output = self.ll_model(base_x)
        if output.dim() > 1 and output.shape[0] > 1:
            output = output.squeeze()
        return loss_fn(output, base_y)

    def step_on_loss(self, loss, optimizer):
        """zero_grad; backward; [data-parallel all-reduce]; clip; step (reference :60-64). With
        FlatAdam the clip is folded into the update kernel."""
        optimizer.zero_grad()
        loss.backward()
        if self.dp is not None:
            self.dp.allreduce_grads(optimizer, self.ll_model)
        if isinstance(optimizer, FlatAdam):
            optimizer.step(max_norm=self.training_args["clip_grad_norm"])
        else:
            self.clip_grad_fn()
            optimizer.step()

    def _use_graphs(self, optimizer) -> bool:
        return (bool(self.training_args.get("cuda_graphs")) and isinstance(optimizer, FlatAdam)
                and not self.training_args["use_single_loss"])

    def _graphed_iit(self, sb, ss, hl_node, loss_fn, optimizer):
        return self._graphed_term(("iit", str(hl_node)), optimizer, lambda: (
            self.get_IIT_loss_over_batch(sb, ss, hl_node, loss_fn) * self.training_args["iit_weight"]))

    def _graphed_behavior(self, sb, loss_fn, optimizer):
        return self._graphed_term(("behavior",), optimizer, lambda: (
            self.get_behaviour_loss_over_batch(sb, loss_fn) * self.training_args["behavior_weight"]))

    def run_train_step(self, base_input, ablation_input, loss_fn, optimizer):
        if self._use_graphs(optimizer):
            sb, ss = self._static_inputs(base_input, ablation_input)
            optimizer.sync_lr()
            hl_node = self.sample_hl_name()
            iit_loss = self._graphed_iit(sb, ss, hl_node, loss_fn, optimizer)
            behavior_loss = self._graphed_behavior(sb, loss_fn, optimizer)
            vals = t.stack([iit_loss, behavior_loss]).tolist()
            return {"train/iit_loss": vals[0], "train/behavior_loss": vals[1]}
        single = self.training_args["use_single_loss"]
        hl_node = self.sample_hl_name()
        iit_loss = (self.get_IIT_loss_over_batch(base_input, ablation_input, hl_node, loss_fn)
                    * self.training_args["iit_weight"])
        if not single:
            self.step_on_loss(iit_loss, optimizer)
        behavior_loss = self.get_behaviour_loss_over_batch(base_input, loss_fn) * self.training_args["behavior_weight"]
        if not single:
            self.step_on_loss(behavior_loss, optimizer)
        else:
            self.step_on_loss(iit_loss + behavior_loss, optimizer)
        vals = t.stack([iit_loss.detach(), behavior_loss.detach()]).tolist()  # one host sync for both
        return {"train/iit_loss": vals[0], "train/behavior_loss": vals[1]}

    def run_eval_step(self, base_input, ablation_input, loss_fn):
        atol = self.training_args["atol"]
        hl_node = self.sample_hl_name()
        hl_output, ll_output = self.do_intervention(base_input, ablation_input, hl_node)
        base_x, base_y, _ = base_input
        loss = loss_fn(ll_output, hl_output)
        if self.hl_model.is_categorical():
            if ll_output.shape == hl_output.shape:  # one-hot labels
                hl_output = ops.argmax_rows(hl_output.reshape(-1, hl_output.shape[-1])).view(hl_output.shape[:-1])
            counts, _ = ops.iia_count(ll_output, hl_output)
            output = self.ll_model(base_x)
            if output.shape == base_y.shape:
                base_y = ops.argmax_rows(base_y.reshape(-1, base_y.shape[-1])).view(base_y.shape[:-1])
            acc_counts, _ = ops.iia_count(output, base_y)
            hits, n, ahits, an = t.cat([counts, acc_counts]).tolist()
            return {"val/iit_loss": loss.item(), "val/IIA": _f32_ratio(hits, n), "val/accuracy": _f32_ratio(ahits, an)}
        IIA = ((ll_output - hl_output).abs() < atol).float().mean().item()
        output = self.ll_model(base_x)
        accuracy = ((output.squeeze() - base_y).abs() < atol).float().mean()
        return {"val/iit_loss": loss.item(), "val/IIA": IIA, "val/accuracy": accuracy.item()}
