"""Plain IIT pair (API mirror of the reference's iit/model_pairs/iit_model_pair.py:6-99)."""
from __future__ import annotations

from typing import Callable

import numpy as np
import torch as t
from torch import Tensor

from iit_b200 import ops
from iit_b200.model_pairs.base_model_pair import BaseModelPair
from iit_b200.utils.metric import MetricStore, MetricStoreCollection, MetricType


def _f32_ratio(num: int, den: int) -> float:
    """What `(a == b).float().mean().item()` returns: a float32 quotient of two exactly representable counts."""
    return float(np.float32(num) / np.float32(den))


class IITModelPair(BaseModelPair):
    def __init__(self, hl_model=None, ll_model=None, hl_graph=None, corr=None, seed=0, training_args=None):
        if hl_model is None:
            assert hl_graph is not None
            hl_model = self.make_hl_model(hl_graph)
        self.hl_model = hl_model
        self.ll_model = ll_model
        self.hl_model.requires_grad_(False)
        self.corr = {} if corr is None else corr
        missing = [str(k) for k in self.corr.keys() if str(k) not in self.hl_model.hook_dict]
        if missing:
            raise AssertionError(f"correspondence keys without an HL hook point: {missing}")
        self.rng = np.random.default_rng(seed)
        defaults = {
            "batch_size": 256,
            "lr": 0.001,
            "num_workers": 0,
            "early_stop": True,
            "lr_scheduler": t.optim.lr_scheduler.ReduceLROnPlateau,
            "scheduler_val_metric": "val/accuracy",
            "scheduler_mode": "max",
            "clip_grad_norm": 1.0,
        }
        self.training_args = {**defaults, **(training_args or {})}
        self.wandb_method = "iit"

    @property
    def loss_fn(self):
        return ops.cross_entropy

    @loss_fn.setter
    def loss_fn(self, value):
        self._loss_fn = value

    @staticmethod
    def make_train_metrics():
        return MetricStoreCollection([MetricStore("train/iit_loss", MetricType.LOSS)])

    @staticmethod
    def make_test_metrics():
        return MetricStoreCollection([MetricStore("val/iit_loss", MetricType.LOSS),
                                      MetricStore("val/accuracy", MetricType.ACCURACY)])

    def run_eval_step(self, base_input, ablation_input, loss_fn: Callable[[Tensor, Tensor], Tensor]):
        hl_node = self.sample_hl_name()
        hl_output, ll_output = self.do_intervention(base_input, ablation_input, hl_node)
        loss = loss_fn(ll_output, hl_output)
        counts, _ = ops.iia_count(ll_output, hl_output)
        hits, n = counts.tolist()
        return {"val/iit_loss": loss.item(), "val/accuracy": _f32_ratio(hits, n)}

    def run_train_step(self, base_input, ablation_input, loss_fn, optimizer):
        optimizer.zero_grad()
        hl_node = self.sample_hl_name()
        loss = self.get_IIT_loss_over_batch(base_input, ablation_input, hl_node, loss_fn)
        loss.backward()
        if self.dp is not None:
            self.dp.allreduce_grads(optimizer, self.ll_model)
        optimizer.step()
        return {"train/iit_loss": loss.item()}
