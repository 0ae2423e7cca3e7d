"""`TorchIndex` / `Ix`: which part of a hooked tensor a node covers.

API mirror of the reference's iit/utils/index.py:9-122 (`as_index`, `hashable_tuple`, hashing,
equality, `intersects`, `Ix[...]` sugar), plus `box()` -- the per-dimension [start, extent)
descriptor the patch kernels consume. Entries may be None (whole dim), int, step-less slice, or a
list of ints (accepted and hashable like the reference; lists have no box form and are routed to
the gather path by the caller).
"""
from __future__ import annotations

import math
from typing import Iterable, Optional


class TorchIndex:
    __slots__ = ("as_index", "hashable_tuple")

    def __init__(self, list_of_things_in_tuple: Iterable):
        items = list(list_of_things_in_tuple)
        for it in items:
            if it is None or type(it) in (int, slice):
                continue
            if not (isinstance(it, list) and all(type(v) is int for v in it)):
                raise AssertionError(f"TorchIndex entries must be None, int, slice or list[int]; got {it!r}")
        self.as_index = tuple(slice(None) if it is None else it for it in items)
        # slices are unhashable before python 3.12: store their reduce form, like the reference does
        self.hashable_tuple = tuple(
            it.__reduce__() if isinstance(it, slice) else (tuple(it) if isinstance(it, list) else it) for it in items)

    def __hash__(self):
        return hash(self.hashable_tuple)

    def __eq__(self, other):
        return isinstance(other, TorchIndex) and self.hashable_tuple == other.hashable_tuple

    def __ne__(self, other):
        return not self.__eq__(other)

    def __repr__(self) -> str:
        parts = []
        for it in self.as_index:
            if isinstance(it, slice):
                if it.step is not None:
                    raise AssertionError("Step is not supported")
                parts.append(":" if it.start is None and it.stop is None
                             else f"{'' if it.start is None else it.start}:{'' if it.stop is None else it.stop}")
            elif isinstance(it, int):
                parts.append(str(it))
            else:
                raise NotImplementedError(it)
        return "[" + ", ".join(parts) + "]"

    def graphviz_index(self, use_actual_colon=True) -> str:
        return repr(self)

    def is_whole(self) -> bool:
        return all(isinstance(it, slice) and it == slice(None) for it in self.as_index)

    def intersects(self, other: Optional["TorchIndex"]) -> bool:
        """Same truth table as the reference (index.py:80-112); None / Ix[[None]] mean everything."""
        if other is None or self == Ix[[None]] or other == Ix[[None]]:
            return True
        if len(self.as_index) != len(other.as_index):
            raise ValueError("Cannot compare indices of different lengths")
        for a, b in zip(self.as_index, other.as_index):
            if a == slice(None) or b == slice(None):
                continue
            a_int, b_int = type(a) is int, type(b) is int
            if a_int and b_int:
                if a != b:
                    return False
            elif not a_int and not b_int:
                lo = 0 if (a.start is None and b.start is None) else max(a.start or 0, b.start or 0)
                hi = min(math.inf if a.stop is None else a.stop, math.inf if b.stop is None else b.stop)
                if lo >= hi:
                    return False
            else:
                sl, i = (b, a) if a_int else (a, b)
                if not ((sl.start or 0) <= i < (math.inf if sl.stop is None else sl.stop)):
                    return False
        return True

    # ---- kernel-side form ------------------------------------------------------------------
    def box(self, shape) -> Optional[list]:
        """[(start, extent)] per tensor dim following basic-indexing rules (missing trailing dims are
        whole; negative ints/bounds wrap; slices clamp). None if a list entry makes it a non-box."""
        idx = self.as_index
        if len(idx) > len(shape):
            raise IndexError(f"too many indices for tensor of dimension {len(shape)}")
        out = []
        for d, size in enumerate(shape):
            it = idx[d] if d < len(idx) else slice(None)
            if isinstance(it, list):
                return None
            if isinstance(it, int):
                i = it + size if it < 0 else it
                if not 0 <= i < size:
                    raise IndexError(f"index {it} is out of bounds for dimension {d} with size {size}")
                out.append((i, 1))
            else:
                if it.step not in (None, 1):
                    return None
                start, stop, _ = it.indices(size)
                out.append((start, max(stop - start, 0)))
        return out


class Index:
    """Syntactic sugar: Ix[:, :, 3] -> TorchIndex((slice(None), slice(None), 3))."""

    def __getitem__(self, index) -> TorchIndex:
        if not isinstance(index, (tuple, list)):
            index = (index,)
        return TorchIndex(index)


Ix = Index()
