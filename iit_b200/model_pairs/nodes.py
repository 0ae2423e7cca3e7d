"""HL / LL node records (API mirror of the reference's iit/model_pairs/nodes.py:10-56)."""
from __future__ import annotations

import dataclasses
from dataclasses import dataclass
from typing import Optional

import torch as t

from iit_b200.utils.index import Ix, TorchIndex

HookName = str
HLCache = dict


@dataclass
class HLNode:
    """Hashes and compares by name only, and equals a plain `str` of the same name, so a
    Correspondence keyed by HLNode can be looked up with either."""
    name: HookName
    num_classes: int
    index: Optional[TorchIndex] = None

    def __post_init__(self):
        if self.index is None:
            self.index = Ix[[None]]

    def __hash__(self) -> int:
        return hash(self.name)

    def __eq__(self, other) -> bool:
        if isinstance(other, HLNode):
            return self.name == other.name
        if isinstance(other, str):
            return self.name == other
        return False

    def __str__(self) -> str:
        return self.name

    __repr__ = __str__


@dataclass
class LLNode:
    name: HookName
    index: Optional[TorchIndex]
    subspace: Optional[t.Tensor] = None

    def __post_init__(self):
        if self.index is None:
            self.index = Ix[[None]]

    def __eq__(self, other):
        return isinstance(other, LLNode) and dataclasses.astuple(self) == dataclasses.astuple(other)

    def __hash__(self):
        return hash(dataclasses.astuple(self))

    def get_index(self):
        return self.index.as_index
