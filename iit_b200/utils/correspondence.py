"""`Correspondence`: HLNode -> set[LLNode], with the hook-name suffixes used to enumerate LL nodes
(API mirror of the reference's iit/utils/correspondence.py:3-70)."""
from __future__ import annotations

from iit_b200.model_pairs.nodes import HLNode, LLNode

_DEFAULT_SUFFIXES = {"attn": "attn.hook_result", "mlp": "mlp.hook_post"}


class Correspondence(dict):
    def __init__(self, *args, suffixes=None, **kwargs):
        super().__init__(*args, **kwargs)
        self.suffixes = dict(_DEFAULT_SUFFIXES) if suffixes is None else suffixes

    def __setattr__(self, key, value):
        if key == "suffixes":
            if not isinstance(value, dict):
                raise AssertionError(f"suffixes must be a dict, got {type(value)}")
        super().__setattr__(key, value)

    def get_suffixes(self):
        return self.suffixes

    @staticmethod
    def get_hook_suffix(corr) -> dict:
        """Everything after 'blocks.<layer>.' for the attn / mlp hooks used by `corr`; at most one
        location per kind (correspondence.py:34-57)."""
        found = {}
        for ll_nodes in corr.values():
            for node in ll_nodes:
                suffix = ".".join(node.name.split(".")[2:])
                kind = "attn" if "attn" in node.name else "mlp" if "mlp" in node.name else None
                if kind is None:
                    raise ValueError(f"Unknown node type {node.name}")
                if found.get(kind, suffix) != suffix:
                    raise ValueError(f"Multiple {kind} suffixes found: {found[kind]} and {suffix}, "
                                     f"multiple {kind} hook locations are not supported yet.")
                found[kind] = suffix
        return found

    @classmethod
    def make_corr_from_dict(cls, d, suffixes=None, make_suffixes_from_corr=False):
        corr = {HLNode(k, -1): {LLNode(name=n, index=None) for n in v} for k, v in d.items()}
        if make_suffixes_from_corr:
            suffixes = cls.get_hook_suffix(corr)
        return cls(corr, suffixes=suffixes)
