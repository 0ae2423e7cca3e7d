"""Enumerating LL nodes inside / outside the circuit (API mirror of the hot-path part of the
reference's iit/utils/node_picker.py:9-54; the parameter pickers are out of scope).

`get_nodes_not_in_circuit` also accepts the ResNet wrapper, for which the reference has no
definition (its StrictIITModelPair needs `cfg.n_heads`): there the candidate nodes are the `conv`
hook points taken whole (iit/utils/wrapper.py:69-70), minus any that the correspondence touches.
"""
from __future__ import annotations

from iit_b200.model_pairs.nodes import LLNode
from iit_b200.utils.index import Ix

_DEFAULT_SUFFIXES = {"attn": "attn.hook_result", "mlp": "mlp.hook_post"}


def get_all_nodes(model, suffixes=None) -> list:
    suffixes = _DEFAULT_SUFFIXES if suffixes is None else suffixes
    if not hasattr(model, "cfg"):  # ResNet-style wrapper
        from iit_b200.utils.wrapper import get_hook_points
        return [LLNode(name, Ix[[None]]) for name in get_hook_points(model)]
    nodes = []
    for layer in range(model.cfg.n_layers):
        attn = f"blocks.{layer}.{suffixes['attn']}"
        nodes.extend(LLNode(attn, Ix[:, :, head, :]) for head in range(model.cfg.n_heads))
        nodes.append(LLNode(f"blocks.{layer}.{suffixes['mlp']}", Ix[[None]]))
    return nodes


def get_nodes_in_circuit(hl_ll_corr) -> list:
    seen = set()
    for ll_nodes in hl_ll_corr.values():
        seen.update(ll_nodes)
    return list(seen)


def nodes_intersect(a: LLNode, b: LLNode) -> bool:
    return a.name == b.name and a.index.intersects(b.index)


def get_nodes_not_in_circuit(ll_model, hl_ll_corr) -> list:
    suffixes = hl_ll_corr.get_suffixes() if hasattr(hl_ll_corr, "get_suffixes") else None
    in_circuit = get_nodes_in_circuit(hl_ll_corr)
    return [n for n in get_all_nodes(ll_model, suffixes) if not any(nodes_intersect(n, c) for c in in_circuit)]
