"""Model-pair base class: the interchange-intervention primitive and the training loop.

API mirror of the reference's iit/model_pairs/base_model_pair.py:23-326 (same method names,
argument meaning, return values, metric bookkeeping), with the device work underneath done by the
sm_100a kernels:
  * `do_intervention` caches ONLY the hook points the Correspondence names (reference :86 caches
    all of them) and runs the source forward without building an autograd graph -- the cache never
    carries gradient in the reference either (transformer_lens stores detached tensors);
  * `make_ll_ablation_hook` returns a hook that calls the patch kernel in place
    (reference :157-161: full clone + strided copy_);
  * HL models that implement `interchange()` evaluate the intervened high-level run in one kernel.
"""
from __future__ import annotations

from abc import ABC, abstractmethod
from typing import Any, Callable

import numpy as np
import torch as t
from torch import Tensor

from iit_b200 import ops
from iit_b200.model_pairs.nodes import HLNode, LLNode
from iit_b200.utils.index import Ix
from iit_b200.utils.metric import MetricStoreCollection, MetricType


class _LazyCache(dict):
    """A cache dict filled by `fill()` the first time anything is read from it."""

    def __init__(self, fill):
        super().__init__()
        self._fill = fill

    def _load(self):
        if self._fill is not None:
            fill, self._fill = self._fill, None
            with t.no_grad():
                self.update(fill())

    def __getitem__(self, k):
        self._load()
        return super().__getitem__(k)

    def __contains__(self, k):
        self._load()
        return super().__contains__(k)

    def __iter__(self):
        self._load()
        return super().__iter__()

    def __len__(self):
        self._load()
        return super().__len__()

    def keys(self):
        self._load()
        return super().keys()

    def items(self):
        self._load()
        return super().items()

    def values(self):
        self._load()
        return super().values()

    def get(self, k, default=None):
        self._load()
        return super().get(k, default)


class BaseModelPair(ABC):
    hl_model: Any
    ll_model: Any
    corr: dict
    training_args: dict
    wandb_method: str
    rng: np.random.Generator
    dp = None  # optional iit_b200.dist.DataParallel

    # ---- to be provided by subclasses (reference :36-69) -------------------------------------
    @property
    @abstractmethod
    def loss_fn(self) -> Callable[[Tensor, Tensor], Tensor]:
        ...

    @staticmethod
    @abstractmethod
    def make_train_metrics() -> MetricStoreCollection:
        ...

    @staticmethod
    @abstractmethod
    def make_test_metrics() -> MetricStoreCollection:
        ...

    @abstractmethod
    def run_train_step(self, base_input, ablation_input, loss_fn, optimizer):
        ...

    @abstractmethod
    def run_eval_step(self, base_input, ablation_input, loss_fn):
        ...

    # ---- intervention ------------------------------------------------------------------------
    def _resolve_hl_node(self, hl_node) -> HLNode:
        if isinstance(hl_node, HLNode):
            return hl_node
        for k in self.corr.keys():  # a str compares equal to the HLNode of that name (nodes.py)
            if k == hl_node:
                return k if isinstance(k, HLNode) else HLNode(str(k), -1)
        raise KeyError(f"{hl_node!r} is not a key of the correspondence")

    def ll_source_run(self, ablation_x, ll_nodes):
        """Source (ablation) forward of the LL model; caches the hook points of `ll_nodes` only.
        Train-mode side effects (BatchNorm running statistics) happen exactly as in a full forward."""
        names = list({n.name for n in ll_nodes})
        # in eval mode nothing after the last cached hook point matters (no BatchNorm statistics move): the sweep
        # sets `stop_source_run_at_hook` and saves that part of the source forward
        # ... and a model whose forward has no side effect in ANY mode (the transformer: no BatchNorm, no dropout) says so
        # with `forward_is_pure`: its source run always stops at the last cached hook point -- for IOI that skips the
        # rest of the blocks and the 16-position unembedding of a run whose logits nobody reads
        stop = ((bool(getattr(self, "stop_source_run_at_hook", False)) and not self.ll_model.training)
                or bool(getattr(self.ll_model, "forward_is_pure", False)))
        with t.no_grad():
            if stop:
                out, cache = self.ll_model.run_with_cache(ablation_x, names_filter=names, stop_when_cached=True)
            else:
                out, cache = self.ll_model.run_with_cache(ablation_x, names_filter=names)
        self.ll_cache = cache
        return out

    def do_intervention(self, base_input, ablation_input, hl_node, verbose=False) -> tuple[Tensor, Tensor]:
        """(hl_output, ll_output) of the base run with `hl_node` / its LL nodes interchanged from the
        ablation (source) run. Reference: base_model_pair.py:75-105."""
        hl_node = self._resolve_hl_node(hl_node)
        ablation_x = ablation_input[0]
        base_x = base_input[0]
        ll_nodes = self.corr[hl_node]
        hl_output = self.hl_intervention(base_input, ablation_input, hl_node)
        self.ll_source_run(ablation_x, ll_nodes)
        # `base_resume` (set by the causality sweep): the base run's prefix up to the hooked block was computed once for all
        # the nodes interchanged at this hook point; the run resumes from that block's cached input
        resume = getattr(self, "base_resume", None)
        kw = {"resume": resume} if resume is not None else {}
        ll_output = self.ll_model.run_with_hooks(
            base_x, fwd_hooks=[(n.name, self.make_ll_ablation_hook(n)) for n in ll_nodes], **kw)
        if verbose:
            print(f"{hl_node=}, {ll_nodes=}\n{hl_output=}")
        return hl_output, ll_output

    def hl_intervention(self, base_input, ablation_input, hl_node: HLNode) -> Tensor:
        """HL source run + HL base run with the node's value swapped (reference :80,89-91)."""
        if hasattr(self.hl_model, "interchange"):
            # the kernel needs no HL cache; callers that read `pair.hl_cache` after an intervention (reference :80) get
            # the source run's cache on first access
            self.hl_cache = _LazyCache(lambda: self.hl_model.run_with_cache(ablation_input)[1])
            return self.hl_model.interchange(base_input, ablation_input, hl_node)
        _, self.hl_cache = self.hl_model.run_with_cache(ablation_input)
        return self.hl_model.run_with_hooks(
            base_input, fwd_hooks=[(hl_node.name, self.make_hl_ablation_hook(hl_node))])

    @staticmethod
    def get_label_idxs():
        return Ix[[None]]

    def make_hl_model(self, hl_graph):
        raise NotImplementedError

    def set_corr(self, corr):
        self.corr = corr

    def sample_hl_name(self) -> HLNode:
        # numpy Generator.choice over the key list: the exact call the reference makes (:117-118), so
        # with the same seed the same nodes are drawn in the same order on every rank.
        return self.rng.choice(list(self.corr.keys()))

    def make_hl_ablation_hook(self, hl_node: HLNode):
        if not isinstance(hl_node, HLNode):
            raise AssertionError(f"hl_node is not an instance of HLNode, but {type(hl_node)}")

        def hl_ablation_hook(hook_point_out: Tensor, hook) -> Tensor:
            cached = self.hl_cache[hook.name]
            if not isinstance(hook_point_out, Tensor):
                return cached
            return ops.patch(hook_point_out, cached, hl_node.index, inplace=False)   # CUDA tensors only, like every operator

        return hl_ablation_hook

    def make_ll_ablation_hook(self, ll_node: LLNode) -> Callable[[Tensor, Any], Tensor]:
        if ll_node.subspace is not None:
            raise NotImplementedError

        def ll_ablation_hook(hook_point_out: Tensor, hook) -> Tensor:
            index = ll_node.index if ll_node.index is not None else Ix[[None]]
            # in place only where the model vouches for it (HookPoint.inplace_ok: nothing saved this tensor for its
            # backward) or no backward will follow; everywhere else clone first, like the reference (:157-161)
            inplace = bool(getattr(hook, "inplace_ok", False)) or not (t.is_grad_enabled() and hook_point_out.requires_grad)
            return ops.patch(hook_point_out, self.ll_cache[hook.name], index, inplace=inplace)

        # (pair, node): lets a layer that can apply the patch itself find its source and index (HookPoint.fusable_patch), and
        # marks the hook as a pure interchange (HookPoint.only_patch_hooks)
        ll_ablation_hook._iit_patch = (self, ll_node)
        return ll_ablation_hook

    def get_IIT_loss_over_batch(self, base_input, ablation_input, hl_node, loss_fn):
        if loss_fn is ops.cross_entropy and hasattr(self.hl_model, "interchange_ce"):
            # the same four runs as do_intervention (reference :75-105, :165-174), with the HL base run and the loss as ONE
            # kernel: the intervened HL output is the class index each row of the LL logits is scored against
            hl_node = self._resolve_hl_node(hl_node)
            ll_nodes = self.corr[hl_node]
            self.hl_cache = _LazyCache(lambda: self.hl_model.run_with_cache(ablation_input)[1])
            self.ll_source_run(ablation_input[0], ll_nodes)
            ll_output = self.ll_model.run_with_hooks(
                base_input[0], fwd_hooks=[(n.name, self.make_ll_ablation_hook(n)) for n in ll_nodes])
            if ll_output.dim() == 2:
                loss, self.last_hl_output = self.hl_model.interchange_ce(base_input, ablation_input, hl_node, ll_output)
                return loss
            return loss_fn(ll_output, self.hl_intervention(base_input, ablation_input, hl_node))
        hl_output, ll_output = self.do_intervention(base_input, ablation_input, hl_node)
        return loss_fn(ll_output, hl_output)

    # ---- CUDA graphs -------------------------------------------------------------------------
    # One loss term of a train step (loss -> zero_grad -> backward -> [all-reduce] -> clip -> Adam) is a fixed
    # sequence of ~300 kernel launches once the interchanged node is chosen. With training_args["cuda_graphs"]
    # each (term, node) variant is captured once into a CUDA graph and replayed afterwards: the host only copies
    # the batch into static buffers, draws the nodes and replays. Requires FlatAdam (device-side step / lr) and
    # fixed batch shapes; all variants share one memory pool.
    def _static_inputs(self, base_input, ablation_input):
        key = tuple((tuple(t.shape), t.dtype) for t in (*base_input, *ablation_input))
        st = getattr(self, "_static", None)
        if st is None or st[0] != key:
            st = (key, tuple(t.clone() for t in base_input), tuple(t.clone() for t in ablation_input))
            self._static = st
            self._graphs = {}
        for dst, src in zip((*st[1], *st[2]), (*base_input, *ablation_input)):
            if dst is not src:
                dst.copy_(src, non_blocking=True)
        refresh = getattr(self.ll_model, "refresh_input_cache", None)
        if refresh is not None:  # per-batch operands derived from the inputs (the stem's im2col): once per step
            refresh(st[1][0])
            refresh(st[2][0])
        return st[1], st[2]

    def _graphed_term(self, key, optimizer, build_loss):
        stale = getattr(self.ll_model, "_filter_shadow_stale", None)
        if stale is not None and stale():    # weights changed outside the optimizer (load_state_dict, manual edits)
            self.ll_model.refresh_filter_shadow()
        graphs = self.__dict__.setdefault("_graphs", {})
        ent = graphs.get(key)
        if ent is None:
            ent = graphs[key] = self._capture_term(optimizer, build_loss)
        ent[0].replay()
        return ent[1]

    def _capture_term(self, optimizer, build_loss):
        if not hasattr(self, "_graph_pool"):
            self._graph_pool = t.cuda.graph_pool_handle()
        buffers = list(self.ll_model.buffers())
        snap_opt = optimizer.state_snapshot()
        snap_buf = [b.clone() for b in buffers]
        cur = t.cuda.current_stream()
        side = t.cuda.Stream()
        side.wait_stream(cur)
        with t.cuda.stream(side):  # eager warm-up: first-launch attribute set-up, autograd worker thread
            self.step_on_loss(build_loss(), optimizer)
        cur.wait_stream(side)
        optimizer.state_restore(snap_opt)  # the warm-up was a real update: undo it
        with t.no_grad():
            for b, s in zip(buffers, snap_buf):
                b.copy_(s)
        for cb in getattr(optimizer, "post_step", ()):   # copies derived from the weights (bf16 filter shadows) follow the
            cb()                                         # restore here, eagerly: the graph then starts from fresh copies
        graph = t.cuda.CUDAGraph()
        with t.cuda.graph(graph, pool=self._graph_pool):
            loss = build_loss()
            self.step_on_loss(loss, optimizer)
            out = loss.detach().clone()
        return graph, out

    # ---- optimisation ------------------------------------------------------------------------
    def clip_grad_fn(self):
        if self.training_args["clip_grad_norm"]:
            t.nn.utils.clip_grad_norm_(self.ll_model.parameters(), self.training_args["clip_grad_norm"])

    def make_optimizer(self):
        """Adam over the LL parameters (reference :229). Flat fused kernel on CUDA."""
        from iit_b200.nn.optim import FlatAdam
        opt = FlatAdam(self.ll_model.parameters(), lr=self.training_args["lr"])
        refresh = getattr(self.ll_model, "refresh_filter_shadow", None)
        if refresh is not None:   # derived copies of the weights follow every update in one launch
            opt.post_step.append(refresh)
        return opt

    def step_scheduler(self, lr_scheduler, test_metrics):
        if isinstance(lr_scheduler, t.optim.lr_scheduler.ReduceLROnPlateau):
            val_metric = self.training_args.get("scheduler_val_metric", "val/accuracy")
            try:
                lr_scheduler.step(test_metrics.to_dict()[val_metric])
            except KeyError:
                raise ValueError(f"val_metric {val_metric} not found in test_metrics {test_metrics}")
            return
        try:
            lr_scheduler.step()
        except Exception as e:  # the reference only warns here
            print(f"WARNING: Could not step lr_scheduler {lr_scheduler} with exception {e}")

    def train(self, train_set, test_set, epochs=1000, use_wandb=False):
        from iit_b200.utils.iit_dataset import IITDataset
        training_args = self.training_args
        print(f"{training_args=}")
        if not isinstance(train_set, IITDataset) or not isinstance(test_set, IITDataset):
            raise AssertionError("train_set and test_set must be IITDataset instances")
        train_loader, test_loader = self.make_loaders(train_set, test_set, training_args["batch_size"],
                                                      training_args["num_workers"])
        early_stop = training_args["early_stop"]
        optimizer = self.make_optimizer()
        loss_fn = self.loss_fn
        scheduler_cls = training_args.get("lr_scheduler", None)
        lr_scheduler = None
        if scheduler_cls == t.optim.lr_scheduler.ReduceLROnPlateau:
            lr_scheduler = scheduler_cls(optimizer, mode=training_args.get("scheduler_mode", "max"), factor=0.1,
                                         patience=10)
        elif scheduler_cls:
            lr_scheduler = scheduler_cls(optimizer)
        wandb = None
        if use_wandb:
            import wandb
            if not wandb.run:
                wandb.init(project="iit")
            wandb.config.update(training_args)
            wandb.config.update({"method": self.wandb_method})
        epoch = -1
        for epoch in range(epochs):
            train_metrics = self._run_train_epoch(train_loader, loss_fn, optimizer)
            test_metrics = self._run_eval_epoch(test_loader, loss_fn)
            if lr_scheduler is not None:
                self.step_scheduler(lr_scheduler, test_metrics)
            self.test_metrics, self.train_metrics = test_metrics, train_metrics
            self._print_and_log_metrics(epoch, train_metrics.metrics + test_metrics.metrics, use_wandb)
            if early_stop and self._check_early_stop_condition(test_metrics.metrics):
                break
        if use_wandb:
            wandb.log({"final epoch": epoch})

    @staticmethod
    def make_loaders(dataset, test_dataset, batch_size, num_workers):
        return dataset.make_loader(batch_size, num_workers), test_dataset.make_loader(batch_size, num_workers)

    def _run_train_epoch(self, loader, loss_fn, optimizer) -> MetricStoreCollection:
        self.ll_model.train()
        train_metrics = self.make_train_metrics()
        for base_input, ablation_input in loader:
            train_metrics.update(self.run_train_step(base_input, ablation_input, loss_fn, optimizer))
        return train_metrics

    def _run_eval_epoch(self, loader, loss_fn) -> MetricStoreCollection:
        self.ll_model.eval()
        test_metrics = self.make_test_metrics()
        with t.no_grad():
            for base_input, ablation_input in loader:
                test_metrics.update(self.run_eval_step(base_input, ablation_input, loss_fn))
        return test_metrics

    @staticmethod
    def _check_early_stop_condition(test_metrics):
        """True once every ACCURACY metric has reached 100 (reference :301-314)."""
        seen = False
        for metric in test_metrics:
            if metric.type == MetricType.ACCURACY:
                seen = True
                if metric.get_value() < 100:
                    return False
        if not seen:
            raise ValueError("No accuracy metric found in test_metrics!")
        return True

    @staticmethod
    def _print_and_log_metrics(epoch, metrics, use_wandb=False):
        print(f"\nEpoch {epoch}:", end=" ")
        if use_wandb:
            import wandb
            wandb.log({"epoch": epoch})
        for metric in metrics:
            print(metric, end=", ")
            if use_wandb:
                import wandb
                wandb.log({metric.get_name(): metric.get_value()})
        print()
