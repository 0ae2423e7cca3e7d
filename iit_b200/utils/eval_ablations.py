"""Resample / mean / zero ablation of LL nodes and their causal effect on the output
(API mirror of the reference's iit/utils/eval_ablations.py:15-295; driver eval_ioi.py:53-71). SURVEY.md 8(f) row 2.

Same functions, arguments and result dictionaries ({LLNode: float}); underneath:
  * the resample source run caches only the ablated node's hook point and, in eval mode, stops there;
  * the ablation hooks are the patch kernel in place (`dst[box] = src[box]` with the per-sample source, the batch
    mean broadcast over the batch, or zeros);
  * KL / argmax / masked counts run in the iit_kl_div_rows and iit_iia_count kernels; LL outputs are unembedded at
    the label position only when the label index is the last position and the model can do it.
Categorical labels may be one-hot ([B, S, V], as the reference's IOI dataset builds them) or class ids ([B, S]).
The pandas / image export helpers of the reference (:262-352) are presentation code and not rebuilt.
"""
from __future__ import annotations

from enum import Enum
from typing import Callable, Dict

import torch as t

from iit_b200 import ops
from iit_b200.model_pairs.nodes import LLNode
from iit_b200.utils.eval_metrics import kl_div
from iit_b200.utils.index import Ix
from iit_b200.utils.node_picker import get_all_nodes, get_nodes_in_circuit, get_nodes_not_in_circuit


class Categorical_Metric(Enum):
    ACCURACY = 1
    KL = 2


def _is_last_pos(label_idx) -> bool:
    return label_idx == Ix[:, -1]


def _ll_forward(model_pair, x, fwd_hooks, label_idx):
    """(output, index to apply to it to reach the label positions)."""
    ll = model_pair.ll_model
    if _is_last_pos(label_idx) and getattr(ll, "supports_last_pos", False):
        return ll.run_with_hooks(x, fwd_hooks=fwd_hooks, last_pos_only=True), Ix[[None]]
    return ll.run_with_hooks(x, fwd_hooks=fwd_hooks), label_idx


def _labels(y: t.Tensor, label_idx) -> t.Tensor:
    """Class ids at the label positions from one-hot (float) or id (int64) labels."""
    ids = y if y.dtype == t.int64 else ops.argmax_rows(y.reshape(-1, y.shape[-1])).view(y.shape[:-1])
    return ids[label_idx.as_index]


def _argmax_at(out: t.Tensor, idx) -> t.Tensor:
    sel = out[idx.as_index]
    return ops.argmax_rows(sel.reshape(-1, sel.shape[-1]).float()).view(sel.shape[:-1])


def do_intervention(model_pair, base_input, ablation_input, node: LLNode, hooker: Callable, label_idx=None):
    """Source run on `ablation_input` (cached at `node` only), base run on `base_input` with `hooker` at the node."""
    ll = model_pair.ll_model
    with t.no_grad():
        if ll.training:
            _, cache = ll.run_with_cache(ablation_input, names_filter=[node.name])
        else:
            _, cache = ll.run_with_cache(ablation_input, names_filter=[node.name], stop_when_cached=True)
    model_pair.ll_cache = cache
    if label_idx is None:
        return ll.run_with_hooks(base_input, fwd_hooks=[(node.name, hooker)])
    return _ll_forward(model_pair, base_input, [(node.name, hooker)], label_idx)


def resample_ablate_node(model_pair, base_in, ablation_in, node: LLNode, results: Dict, hooker: Callable, atol=5e-2,
                         verbose=False, categorical_metric: Categorical_Metric = Categorical_Metric.KL):
    """Adds this batch's effect of replacing `node` by its value on the ablation input (eval_ablations.py:35-125)."""
    base_x, base_y, _ = base_in
    ablation_x, ablation_y, _ = ablation_in
    base_hl_out = model_pair.hl_model(base_in).squeeze()
    if model_pair.hl_model.is_categorical():
        label_idx = model_pair.get_label_idxs()
        ll_out, ll_idx = do_intervention(model_pair, base_x, ablation_x, node, hooker, label_idx)
        label_unchanged = _labels(base_y, label_idx) == _labels(ablation_y, label_idx)
        if categorical_metric == Categorical_Metric.KL:
            hl_sel = base_hl_out[label_idx.as_index]
            ll_sel = ll_out[ll_idx.as_index]
            results[node] += kl_div(ll_sel, hl_sel, Ix[[None]]).mean().item()
        else:
            ll_top = _argmax_at(ll_out, ll_idx)
            hl_top = _argmax_at(base_hl_out, label_idx)
            changed = (~label_unchanged) & (ll_top != hl_top)
            num, den = t.stack([changed.sum(), (~label_unchanged).sum()]).tolist()
            results[node] += float(num) / float(den)
        return
    ll_out = do_intervention(model_pair, base_x, ablation_x, node, hooker)
    label_unchanged = base_y == ablation_y
    ll_unchanged = t.isclose(ll_out.float().squeeze(), base_hl_out.float().to(ll_out.device), atol=atol)
    changed = (~label_unchanged).float() * (~ll_unchanged).float()
    results[node] += changed.sum().item() / (~label_unchanged).float().sum().item()


def _pick_nodes(model_pair, node_type: str, with_suffixes: bool):
    assert node_type in ["a", "c", "n"], "type must be one of 'a', 'c', or 'n'"
    if node_type == "n":
        return get_nodes_not_in_circuit(model_pair.ll_model, model_pair.corr)
    if node_type == "a":
        suffixes = model_pair.corr.get_suffixes() if (with_suffixes and hasattr(model_pair.corr, "get_suffixes")) else None
        return get_all_nodes(model_pair.ll_model, suffixes)
    return get_nodes_in_circuit(model_pair.corr)


def check_causal_effect(model_pair, dataset, batch_size: int = 256, node_type: str = "a", verbose: bool = False,
                        categorical_metric: Categorical_Metric = Categorical_Metric.KL):
    """Mean over batches of the resample-ablation effect per node (eval_ablations.py:128-161)."""
    all_nodes = _pick_nodes(model_pair, node_type, with_suffixes=False)
    hookers = {node: model_pair.make_ll_ablation_hook(node) for node in all_nodes}
    results = {node: 0 for node in all_nodes}
    loader = dataset.make_loader(batch_size=batch_size, num_workers=0)
    with t.no_grad():
        for base_in, ablation_in in loader:
            for node, hooker in hookers.items():
                resample_ablate_node(model_pair, base_in, ablation_in, node, results, hooker, verbose=verbose,
                                     categorical_metric=categorical_metric)
    return {node: result / len(loader) for node, result in results.items()}


def get_mean_cache(model_pair, dataset, batch_size=8, names=None):
    """{hook point: mean activation over the dataset, shape [1, ...]} (eval_ablations.py:164-173). `names` limits the
    hook points (default: every hook point this LL model can expose)."""
    ll = model_pair.ll_model
    if names is None:
        unsupported = set(getattr(ll, "_unsupported", ()))
        names = [n for n in ll.hook_dict if n not in unsupported]
    loader = dataset.make_loader(batch_size=batch_size, num_workers=0)
    mean_cache = {}
    with t.no_grad():
        for batch in loader:
            _, cache = ll.run_with_cache(batch[0], names_filter=list(names))
            for node, tensor in cache.items():
                if node not in mean_cache:
                    mean_cache[node] = t.zeros_like(tensor[0].unsqueeze(0), dtype=t.float32)
                mean_cache[node] += tensor.float().mean(dim=0).unsqueeze(0) / len(loader)
    return mean_cache


def make_ablation_hook(node: LLNode, mean_cache, use_mean_cache: bool = True) -> Callable:
    """In-place hook that overwrites the node's slice with zeros or with the dataset mean (eval_ablations.py:176-193)."""
    if node.subspace is not None:
        raise NotImplementedError("Subspace not supported yet.")
    index = node.index if node.index is not None else Ix[[None]]

    def zero_hook(hook_point_out, hook):
        return ops.zero_box_(hook_point_out, index)

    def mean_hook(hook_point_out, hook):
        cached = mean_cache[node.name].to(hook_point_out.dtype).expand_as(hook_point_out)   # batch stride 0
        return ops.patch_(hook_point_out, cached, index)

    return mean_hook if use_mean_cache else zero_hook


def ablate_node(model_pair, base_in, node: LLNode, results: Dict, hook: Callable, atol=5e-2, verbose=False):
    """Adds, for this batch, the fraction of correctly-predicted samples whose prediction the ablation changes
    (eval_ablations.py:196-237)."""
    base_x, base_y, _ = base_in
    base_hl_out = model_pair.hl_model(base_in).squeeze()
    if model_pair.hl_model.is_categorical():
        label_idx = model_pair.get_label_idxs()
        ll_out, ll_idx = _ll_forward(model_pair, base_x, [(node.name, hook)], label_idx)
        base_ll_out, _ = _ll_forward(model_pair, base_x, [], label_idx)
        ll_top, base_ll_top = _argmax_at(ll_out, ll_idx), _argmax_at(base_ll_out, ll_idx)
        hl_top = _argmax_at(base_hl_out, label_idx)
        accuracy = base_ll_top == hl_top
        changed = (ll_top != hl_top) & accuracy
        num, den = t.stack([changed.sum(), accuracy.sum()]).tolist()
        results[node] += float(num) / (float(den) + 1e-6)
        return
    ll_out = model_pair.ll_model.run_with_hooks(base_x, fwd_hooks=[(node.name, hook)])
    base_ll_out = model_pair.ll_model(base_x).squeeze()
    ll_unchanged = t.isclose(ll_out.float().squeeze(), base_hl_out.float().to(ll_out.device), atol=atol)
    accuracy = t.isclose(base_ll_out.float(), base_hl_out.float(), atol=atol).float()
    changed = (~ll_unchanged).float() * accuracy
    results[node] += changed.sum().item() / (accuracy.sum().item() + 1e-6)


def check_causal_effect_on_ablation(model_pair, dataset, batch_size: int = 256, node_type: str = "a", mean_cache=None,
                                    verbose: bool = False):
    """Mean over batches of the mean- (or zero-) ablation effect per node (eval_ablations.py:262-295)."""
    use_mean_cache = True if mean_cache else False
    all_nodes = _pick_nodes(model_pair, node_type, with_suffixes=True)
    hookers = {node: make_ablation_hook(node, mean_cache, use_mean_cache) for node in all_nodes}
    results = {node: 0 for node in all_nodes}
    loader = dataset.make_loader(batch_size=batch_size, num_workers=0)
    with t.no_grad():
        for base_in in loader:
            for node, hooker in hookers.items():
                ablate_node(model_pair, base_in, node, results, hooker, verbose=verbose)
    return {node: result / len(loader) for node, result in results.items()}


def get_causal_effects_for_all_nodes(model_pair, uni_test_set, batch_size=256, use_mean_cache=True):
    """(effects of nodes not in the circuit, effects of nodes in the circuit) (eval_ablations.py:240-259)."""
    mean_cache = get_mean_cache(model_pair, uni_test_set, batch_size=batch_size) if use_mean_cache else None
    not_in = check_causal_effect_on_ablation(model_pair, uni_test_set, node_type="n", verbose=False, mean_cache=mean_cache)
    in_c = check_causal_effect_on_ablation(model_pair, uni_test_set, node_type="c", verbose=False, mean_cache=mean_cache)
    return not_in, in_c
