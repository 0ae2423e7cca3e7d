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


# ---- parameter pickers (reference: iit/utils/node_picker.py:57-170), used by the freeze / stop-gradient pairs --------
def get_post_nodes_not_in_circuit(ll_model, hl_ll_corr) -> list:
    """One whole-tensor node per hook point that holds a node outside the circuit (reference :57-79)."""
    suffixes = hl_ll_corr.get_suffixes()
    post_nodes = []
    for node in get_nodes_not_in_circuit(ll_model, hl_ll_corr):
        layer = int(node.name.split(".")[1])
        if "attn" in node.name and "attn" in suffixes:
            name = f"blocks.{layer}.{suffixes['attn']}"
        elif "mlp" in suffixes:
            name = f"blocks.{layer}.{suffixes['mlp']}"
        else:
            raise KeyError(f"no hook suffix for {node.name}")
        if all(pn.name != name for pn in post_nodes):
            post_nodes.append(LLNode(name, Ix[[None]]))
    return post_nodes


def _get_param_idx(name: str, param, node: LLNode):
    """Index of the slice of parameter `name` that belongs to `node` (reference :82-111)."""
    from iit_b200.utils.index import TorchIndex
    param_type = name.split(".")[-1]
    node_idx, none_ix = node.index, Ix[[None]]
    if node.subspace is not None:
        raise NotImplementedError("Subspaces are not supported")
    if node_idx == none_ix or param_type == "b_O":
        param_idx = none_ix
    elif param_type in ("W_Q", "W_K", "W_V"):
        param_idx = TorchIndex(node_idx.as_index[:-1])
    elif param_type == "W_O":
        i = list(node_idx.as_index[:-1])
        param_idx = TorchIndex([i[0], i[2], i[1]])
    elif param_type in ("b_Q", "b_K", "b_V"):
        i = list(node_idx.as_index[:-1])
        param_idx = TorchIndex([slice(None), i[2]])
    else:
        raise NotImplementedError(f"Param of type '{param_type}' is expected to have index {none_ix}, but got {node_idx}")
    try:
        param[param_idx.as_index]
    except IndexError:
        raise IndexError(f"Index {param_idx} is out of bounds for param {name}")
    return param_idx


def get_activation_idx(node: LLNode):
    """reference :114-125"""
    from iit_b200.utils.index import TorchIndex
    param_type = node.name.split(".")[-1]
    i = node.index.as_index
    if param_type in ("W_Q", "W_K", "W_V"):
        return TorchIndex([slice(None), *i])
    if param_type in ("b_Q", "b_K", "b_V"):
        return TorchIndex([slice(None), *i, slice(None)])
    if param_type == "W_O":
        return TorchIndex([i[0], i[2], i[1], slice(None)])
    return Ix[[None]]


def get_params_in_circuit(hl_ll_corr, ll_model) -> list:
    """reference :128-140"""
    out = []
    for name, param in ll_model.named_parameters():
        for node in get_nodes_in_circuit(hl_ll_corr):
            if ".".join(node.name.split(".")[:-1]) == ".".join(name.split(".")[:-1]):
                out.append(LLNode(name, _get_param_idx(name, param, node)))
    return out


def get_all_params(ll_model) -> list:
    """reference :143-154: per-head slices of the attention parameters, every other parameter whole."""
    params = []
    for name, param in ll_model.named_parameters():
        if name.split(".")[-1] in ("W_Q", "W_K", "W_V", "W_O", "b_Q", "b_K", "b_V"):
            for head in range(ll_model.cfg.n_heads):
                params.append(LLNode(name, _get_param_idx(name, param, LLNode(name, Ix[:, :, head, :]))))
        else:
            params.append(LLNode(name, Ix[[None]]))
    return params


def get_params_not_in_circuit(hl_ll_corr, ll_model, filter_out_embed: bool = True) -> list:
    """reference :157-170. As written it compares PARAMETER names with HOOK-POINT names (`nodes_intersect` wants equal
    names), which never match: every non-embedding parameter slice comes back as "not in the circuit"."""
    in_circuit = get_nodes_in_circuit(hl_ll_corr)
    return [p for p in get_all_params(ll_model)
            if not (filter_out_embed and "embed" in p.name) and not any(nodes_intersect(p, c) for c in in_circuit)]
