"""Hook points for the low- and high-level models.

The reference builds on transformer_lens' `HookPoint` / `HookedRootModule`
(iit/model_pairs/base_model_pair.py:6-7,80-98; iit/utils/wrapper.py:3-33). This is an independent,
minimal implementation of that interface (same method names, hook signature `hook(tensor, hook=hp)`,
return value replaces the tensor) so the model-pair code and user hooks read the same. Unlike
transformer_lens it does not go through `nn.Module` forward hooks: a HookPoint with no hooks is a
plain pass-through, and models ask `has_hooks()` to decide whether two kernels may be fused across it.
"""
from __future__ import annotations

from contextlib import contextmanager
from typing import Callable, Optional, Sequence, Union

import torch
import torch.nn as nn

NamesFilter = Optional[Union[Callable[[str], bool], Sequence[str], str]]


class HookPoint(nn.Module):
    def __init__(self, inplace_ok: bool = False):
        """`inplace_ok`: the model promises that no operator saved the tensor passing through here for its backward, so
        a patch hook may overwrite it in place. False (the default) makes the patch clone first, as the reference always
        does (iit/model_pairs/base_model_pair.py:157): e.g. a ReLU output is kept by its producer for the backward, and
        overwriting it would trip autograd's version check."""
        super().__init__()
        self.inplace_ok = bool(inplace_ok)
        self.name: Optional[str] = None
        self._fwd: list = []   # (fn, is_permanent, level)
        self._bwd: list = []
        self.ctx = {}

    def add_hook(self, hook, dir="fwd", is_permanent=False, level=None, prepend=False):
        if dir not in ("fwd", "bwd"):
            raise ValueError(f"Invalid direction {dir}")
        lst = self._fwd if dir == "fwd" else self._bwd
        entry = (hook, is_permanent, level)
        lst.insert(0, entry) if prepend else lst.append(entry)

    def remove_hooks(self, dir="fwd", including_permanent=False, level=None):
        def keep(e):
            _, perm, lvl = e
            return not (including_permanent or (not perm and (level is None or lvl == level)))
        if dir in ("fwd", "both"):
            self._fwd = [e for e in self._fwd if keep(e)]
        if dir in ("bwd", "both"):
            self._bwd = [e for e in self._bwd if keep(e)]

    def has_hooks(self) -> bool:
        return bool(self._fwd or self._bwd)

    def only_patch_hooks(self) -> bool:
        """True when nothing but interchange patches (base_model_pair.make_ll_ablation_hook marks them `_iit_patch`) is
        attached: their effect on the gradient is to zero a box, which commutes with an element-wise factor, so a model may
        keep a fused backward across this hook point."""
        return not self._bwd and all(getattr(fn, "_iit_patch", False) for fn, _, _ in self._fwd)

    def fusable_patch(self):
        """(cached source activation, TorchIndex) when the ONLY thing attached here is one interchange patch of a model pair
        (base_model_pair.make_ll_ablation_hook) whose source run has cached this hook point -- the producing layer may then
        apply the patch itself and skip the hook call; None otherwise."""
        if self._bwd or len(self._fwd) != 1:
            return None
        tag = getattr(self._fwd[0][0], "_iit_patch", None)
        if not isinstance(tag, tuple):
            return None
        pair, node = tag
        cache = getattr(pair, "ll_cache", None)
        if node.subspace is not None or cache is None or self.name not in cache:
            return None
        return cache[self.name], node.index

    def clear_context(self):
        self.ctx = {}

    def forward(self, x):
        for fn, _, _ in self._fwd:
            out = fn(x, hook=self)
            if out is not None:
                x = out
        if self._bwd and isinstance(x, torch.Tensor) and x.requires_grad:
            for fn, _, _ in self._bwd:
                x.register_hook(lambda g, fn=fn: fn(g, hook=self))
        return x

    def layer(self) -> int:
        return int(self.name.split(".")[1])


def _as_filter(names_filter: NamesFilter):
    if names_filter is None:
        return lambda name: True
    if isinstance(names_filter, str):
        return lambda name: name == names_filter
    if callable(names_filter):
        return names_filter
    wanted = set(names_filter)
    return lambda name: name in wanted


class _StopForward(Exception):
    """Raised by run_with_cache(stop_when_cached=True) from the last wanted hook point."""


class HookedRootModule(nn.Module):
    """Root of a hooked model. Call `setup()` at the end of __init__."""

    def __init__(self, *args):
        super().__init__()
        self.is_caching = False
        self.context_level = 0

    def setup(self):
        self.mod_dict = {}
        self.hook_dict = {}
        for name, module in self.named_modules():
            if name == "":
                continue
            module.name = name
            self.mod_dict[name] = module
            if isinstance(module, HookPoint):
                self.hook_dict[name] = module

    def hook_points(self):
        return self.hook_dict.values()

    def reset_hooks(self, clear_contexts=True, direction="both", including_permanent=False, level=None):
        for hp in self.hook_points():
            if clear_contexts:
                hp.clear_context()
            hp.remove_hooks(direction, including_permanent=including_permanent, level=level)
        self.is_caching = False

    def add_hook(self, name, hook, dir="fwd", is_permanent=False, level=None, prepend=False):
        if isinstance(name, str):
            if name not in self.hook_dict:
                raise KeyError(f"no hook point named {name!r}")
            self.hook_dict[name].add_hook(hook, dir=dir, is_permanent=is_permanent, level=level, prepend=prepend)
        else:
            for hp_name, hp in self.hook_dict.items():
                if name(hp_name):
                    hp.add_hook(hook, dir=dir, is_permanent=is_permanent, level=level, prepend=prepend)

    @contextmanager
    def hooks(self, fwd_hooks=(), bwd_hooks=(), reset_hooks_end=True, clear_contexts=False):
        self.context_level += 1
        try:
            for name, hook in fwd_hooks:
                self.add_hook(name, hook, dir="fwd", level=self.context_level)
            for name, hook in bwd_hooks:
                self.add_hook(name, hook, dir="bwd", level=self.context_level)
            yield self
        finally:
            if reset_hooks_end:
                self.reset_hooks(clear_contexts, including_permanent=False, level=self.context_level)
            self.context_level -= 1

    def run_with_hooks(self, *model_args, fwd_hooks=(), bwd_hooks=(), reset_hooks_end=True, clear_contexts=False,
                       **model_kwargs):
        with self.hooks(fwd_hooks, bwd_hooks, reset_hooks_end, clear_contexts) as m:
            return m.forward(*model_args, **model_kwargs)

    def get_caching_hooks(self, names_filter: NamesFilter = None, incl_bwd=False, device=None,
                          remove_batch_dim=False, cache=None):
        cache = {} if cache is None else cache
        flt = _as_filter(names_filter)

        def save_hook(tensor, hook):
            v = tensor.detach()  # the cache never carries gradient (base_model_pair.py:86 semantics)
            if device is not None:
                v = v.to(device)
            cache[hook.name] = v[0] if remove_batch_dim else v

        return cache, [(n, save_hook) for n in self.hook_dict if flt(n)], []

    def run_with_cache(self, *model_args, names_filter: NamesFilter = None, device=None, remove_batch_dim=False,
                       incl_bwd=False, reset_hooks_end=True, clear_contexts=False, stop_when_cached=False, **model_kwargs):
        """`stop_when_cached` (not in transformer_lens): with an explicit list of names, abandon the forward pass once
        every listed hook point has been cached; the returned output is then None. Only meaningful when the rest of
        the pass has no side effect the caller wants (eval mode: no BatchNorm statistics to update)."""
        cache, fwd, bwd = self.get_caching_hooks(names_filter, incl_bwd, device, remove_batch_dim)
        if stop_when_cached:
            if names_filter is None or callable(names_filter):
                raise ValueError("stop_when_cached needs an explicit list of hook-point names")
            wanted = {names_filter} if isinstance(names_filter, str) else set(names_filter)
            # A hook point that fires several times per forward (a BasicBlock's shared ReLU) ends up cached at its LAST
            # firing, in transformer_lens and here; stopping at the first would keep the wrong activation. The model
            # lists such hook points in `multi_fire_hook_points`: wanting one of them means running the whole forward.
            if not wanted & set(getattr(self, "multi_fire_hook_points", ())):

                def stop_hook(tensor, hook):
                    if wanted.issubset(cache.keys()):
                        raise _StopForward()

                fwd = fwd + [(n, stop_hook) for n in wanted]
        with self.hooks(fwd_hooks=fwd, bwd_hooks=bwd, reset_hooks_end=reset_hooks_end, clear_contexts=clear_contexts):
            try:
                out = self(*model_args, **model_kwargs)
            except _StopForward:
                out = None
        return out, cache
