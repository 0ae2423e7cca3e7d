"""TEST INFRASTRUCTURE (oracle) -- restatement of transformer_lens==1.12.1 `hook_points`.

transformer-lens is an un-vendored third-party dependency of the reference
(/root/reference/pyproject.toml:11, poetry.lock:3154-3155) and is absent from this image.
This file restates the published behaviour of its hook API from knowledge of 1.12.1
[TL-knowledge]; it exists only so the reference's own Python (`/root/reference/iit/**`)
can be imported and run as the parity oracle. It is never imported by the product path.

Reference call sites this must serve: iit/model_pairs/base_model_pair.py:80,86,89,92;
iit/utils/wrapper.py:3-33; iit/tasks/mnist_pvr/pvr_hl.py:17-24; iit/tasks/ioi/ioi_hl.py:75-97;
iit/model_pairs/strict_iit_model_pair.py:63-67; iit/model_pairs/stop_grad_pair.py:63-93.
"""
from __future__ import annotations

from contextlib import contextmanager
from typing import Callable, Optional, Sequence, Union

import torch
import torch.nn as nn


class LensHandle:
    def __init__(self, hook, is_permanent=False, context_level=None):
        self.hook = hook
        self.is_permanent = is_permanent
        self.context_level = context_level


class HookPoint(nn.Module):
    """Identity module; hooks receive (tensor, hook=self) and may return a replacement."""

    def __init__(self):
        super().__init__()
        self.fwd_hooks: list[LensHandle] = []
        self.bwd_hooks: list[LensHandle] = []
        self.ctx = {}
        self.name: Optional[str] = None

    def add_hook(self, hook, dir="fwd", is_permanent=False, level=None, prepend=False):
        if dir == "fwd":
            def full_hook(module, module_input, module_output):
                return hook(module_output, hook=self)

            handle = self.register_forward_hook(full_hook)
            self.fwd_hooks.append(LensHandle(handle, is_permanent, level))
        elif dir == "bwd":
            def full_hook(module, grad_input, grad_output):
                return hook(grad_output[0], hook=self)

            handle = self.register_full_backward_hook(full_hook)
            self.bwd_hooks.append(LensHandle(handle, is_permanent, level))
        else:
            raise ValueError(f"Invalid direction {dir}")

    def remove_hooks(self, dir="fwd", including_permanent=False, level=None):
        def _remove(handles):
            keep = []
            for h in handles:
                if including_permanent or (not h.is_permanent and (level is None or h.context_level == level)):
                    h.hook.remove()
                else:
                    keep.append(h)
            return keep

        if dir in ("fwd", "both"):
            self.fwd_hooks = _remove(self.fwd_hooks)
        if dir in ("bwd", "both"):
            self.bwd_hooks = _remove(self.bwd_hooks)

    def clear_context(self):
        self.ctx = {}

    def forward(self, x):
        return x

    def layer(self):
        return int(self.name.split(".")[1])


NamesFilter = Optional[Union[Callable[[str], bool], Sequence[str], str]]


class HookedRootModule(nn.Module):
    def __init__(self, *args):
        super().__init__()
        self.is_caching = False
        self.context_level = 0

    def setup(self):
        self.mod_dict = {}
        self.hook_dict: dict[str, HookPoint] = {}
        for name, module in self.named_modules():
            if name == "":
                continue
            module.name = name
            self.mod_dict[name] = module
            if "HookPoint" in str(type(module)):
                self.hook_dict[name] = module

    def hook_points(self):
        return self.hook_dict.values()

    def remove_all_hook_fns(self, direction="both", including_permanent=False, level=None):
        for hp in self.hook_points():
            hp.remove_hooks(direction, including_permanent=including_permanent, level=level)

    def clear_contexts(self):
        for hp in self.hook_points():
            hp.clear_context()

    def reset_hooks(self, clear_contexts=True, direction="both", including_permanent=False, level=None):
        if clear_contexts:
            self.clear_contexts()
        self.remove_all_hook_fns(direction, including_permanent, level=level)
        self.is_caching = False

    def check_and_add_hook(self, hook_point, hook_point_name, hook, dir="fwd", is_permanent=False,
                           level=None, prepend=False):
        hook_point.add_hook(hook, dir=dir, is_permanent=is_permanent, level=level, prepend=prepend)

    def add_hook(self, name, hook, dir="fwd", is_permanent=False, level=None, prepend=False):
        if isinstance(name, str):
            self.check_and_add_hook(self.mod_dict[name], name, hook, dir, is_permanent, level, prepend)
        else:
            for hook_point_name, hp in self.hook_dict.items():
                if name(hook_point_name):
                    self.check_and_add_hook(hp, hook_point_name, hook, dir, is_permanent, level, prepend)

    @contextmanager
    def hooks(self, fwd_hooks=[], bwd_hooks=[], reset_hooks_end=True, clear_contexts=False):
        try:
            self.context_level += 1
            for name, hook in fwd_hooks:
                self.add_hook(name, hook, dir="fwd", level=self.context_level)
            for name, hook in bwd_hooks:
                self.add_hook(name, hook, dir="bwd", level=self.context_level)
            yield self
        finally:
            if reset_hooks_end:
                self.reset_hooks(clear_contexts, including_permanent=False, level=self.context_level)
            self.context_level -= 1

    def run_with_hooks(self, *model_args, fwd_hooks=[], bwd_hooks=[], reset_hooks_end=True,
                       clear_contexts=False, **model_kwargs):
        with self.hooks(fwd_hooks, bwd_hooks, reset_hooks_end, clear_contexts) as hooked_model:
            return hooked_model.forward(*model_args, **model_kwargs)

    def get_caching_hooks(self, names_filter: NamesFilter = None, incl_bwd=False, device=None,
                          remove_batch_dim=False, cache=None):
        if cache is None:
            cache = {}
        if names_filter is None:
            names_filter = lambda name: True
        elif isinstance(names_filter, str):
            filter_str = names_filter
            names_filter = lambda name: name == filter_str
        elif isinstance(names_filter, (list, tuple)):
            filter_list = list(names_filter)
            names_filter = lambda name: name in filter_list

        def save_hook(tensor, hook):
            # [TL-knowledge] the cache stores detached tensors: no gradient flows through a cached run.
            if remove_batch_dim:
                cache[hook.name] = tensor.detach().to(device)[0]
            else:
                cache[hook.name] = tensor.detach().to(device)

        fwd_hooks = [(name, save_hook) for name in self.hook_dict if names_filter(name)]
        return cache, fwd_hooks, []

    def run_with_cache(self, *model_args, names_filter: NamesFilter = None, device=None,
                       remove_batch_dim=False, incl_bwd=False, reset_hooks_end=True,
                       clear_contexts=False, **model_kwargs):
        cache_dict, fwd, bwd = self.get_caching_hooks(names_filter, incl_bwd, device, remove_batch_dim)
        with self.hooks(fwd_hooks=fwd, bwd_hooks=bwd, reset_hooks_end=reset_hooks_end,
                        clear_contexts=clear_contexts):
            model_out = self(*model_args, **model_kwargs)
        return model_out, cache_dict
