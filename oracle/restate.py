"""TEST INFRASTRUCTURE (oracle) -- CPU restatement of the reference's interchange-intervention path.

This module is the CHECKER for the CUDA path. It is never imported by `iit_b200` (the product);
only `tests/`, `__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs may
use it. It restates, in plain torch-CPU fp32 / int64 eager arithmetic, what the reference computes
on the hot path; every function cites the reference file:line it follows (paths relative to
/root/reference). It travels to the GPU box (the reference tree does not).

PINNING: tests/test_oracle_golden.py checks every function here against
  * the reference's own golden vectors (iit/tasks/ioi/test_ioi.py:10-45, tests/test_index.py,
    tests/test_node_picker.py:6-29), copied as literals into the tests, and
  * tests/golden/*.npz, produced by oracle/make_golden.py by RUNNING THE UNMODIFIED REFERENCE
    (/root/reference/iit/**) in the build container on seeded inputs, through the transformer_lens
    shim in oracle/tl_shim (transformer_lens==1.12.1 is an absent third-party dependency; the shim's
    fidelity is [TL-knowledge]).
"""
from __future__ import annotations

import math
from typing import Callable, Optional

import numpy as np
import torch
import torch.nn.functional as F

MNIST_CLASS_MAP = [1, 1, 1, 1, 2, 2, 2, 3, 3, 3]  # iit/tasks/mnist_pvr/utils.py:10


# --------------------------------------------------------------------------------------------
# L2: index / patch  (iit/utils/index.py:23-47, iit/model_pairs/base_model_pair.py:151-163)
# --------------------------------------------------------------------------------------------
def as_index(spec):
    """iit/utils/index.py:36-38 -- None -> slice(None); ints, slices and int lists pass through."""
    if spec is None:
        return (slice(None),)
    return tuple(slice(None) if s is None else s for s in spec)


def patch(act: torch.Tensor, cached: torch.Tensor, spec) -> torch.Tensor:
    """base_model_pair.py:157-161 -- out = act.clone(); out[idx] = cache[name][idx]."""
    out = act.clone()
    idx = as_index(spec)
    out[idx] = cached[idx]
    return out


# --------------------------------------------------------------------------------------------
# HL models
# --------------------------------------------------------------------------------------------
PVR_HOOKS = ["hook_tl", "hook_tr", "hook_bl", "hook_br"]


def pvr_hl_forward(vars_: torch.Tensor) -> torch.Tensor:
    """iit/tasks/mnist_pvr/pvr_hl.py:44-56 -- pointer = class_map[tl]-1; out = stack(tr,bl,br)[pointer, arange]."""
    class_map = torch.tensor(MNIST_CLASS_MAP, dtype=torch.long)
    tl, tr, bl, br = [vars_[:, i] for i in range(4)]
    pointer = class_map[tl] - 1
    stacked = torch.stack([tr, bl, br], dim=0)
    return stacked[pointer, torch.arange(len(pointer))]


def pvr_var_index(hook_name: str) -> int:
    """pvr_hl.py:30-42 / pvr_check_leaky_hl.py:33-43 -- which intermediate var a hook carries.
    Leaky names `hook_{i}_leaked_to_{j}` act on variable i (substring test on 'hook_{i}')."""
    for i, h in enumerate(PVR_HOOKS):
        if h in hook_name:
            return i
    raise ValueError(hook_name)


def pvr_hl_intervened(base_vars, src_vars, hl_node_name: str) -> torch.Tensor:
    """base_model_pair.py:80,89-91 + 120-142 for the PVR HL: the hooked variable of the base run is
    replaced wholesale (index Ix[[None]], pvr_hl.py:60-65) by the source run's cached value."""
    v = base_vars.clone()
    i = pvr_var_index(hl_node_name)
    v[:, i] = src_vars[:, i]
    return pvr_hl_forward(v)


def leaky_target_quadrant(hl_node_name: str) -> int:
    """pvr_check_leaky_hl.py:88-107 -- 'to_xx' picks the LL quadrant that is patched."""
    for i, q in enumerate(["tl", "tr", "bl", "br"]):
        if f"to_{q}" in hl_node_name:
            return i
    raise ValueError(hl_node_name)


def ioi_duplicate_head(tokens: torch.Tensor) -> torch.Tensor:
    """iit/tasks/ioi/ioi_hl.py:16-24 -- last earlier position holding the same token, else -1.
    (CPU index_put with duplicate indices is last-write-wins in nonzero() order == max earlier
    position; restated directly as that max.)"""
    b, s = tokens.shape
    out = torch.full_like(tokens, -1)
    for j in range(s):
        for i in range(j):
            same = tokens[:, i] == tokens[:, j]
            out[:, j] = torch.where(same, torch.full_like(out[:, j], i), out[:, j])
    return out


def ioi_s_inhibition_head(tokens, duplicate):
    """ioi_hl.py:37-46 -- token where a duplicate exists, else -1."""
    ret = tokens.clone()
    ret[duplicate == -1] = -1
    return ret


def ioi_name_mover_head(tokens, s_inhibition, names, d_vocab):
    """ioi_hl.py:54-68 -- +10 at tok if tok in names; += -15 at s_inh where != -1 (index -1 wraps to
    the last vocab slot and adds 0); cumsum over sequence."""
    b, s = tokens.shape
    logits = torch.zeros((b, s, d_vocab))
    name_mask = torch.isin(tokens, names)
    bi, si = torch.meshgrid(torch.arange(b), torch.arange(s), indexing="ij")
    logits[bi, si, tokens] = 10 * name_mask.float()
    logits[bi, si, s_inhibition] += -15 * s_inhibition.ne(-1).float()
    return torch.cumsum(logits, dim=1)


IOI_HL_HOOKS = ["all_nodes_hook", "hook_duplicate", "hook_s_inhibition", "hook_name_mover"]


def ioi_hl_forward(tokens, names, d_vocab, replace: Optional[dict] = None, cache: Optional[dict] = None):
    """ioi_hl.py:102-130. `replace[name]` substitutes the value at that hook point (whole-tensor
    interchange, correspondence.py:66 index=None); `cache` collects hook-point values."""
    replace = replace or {}

    def hp(name, val):
        if name in replace:
            val = replace[name]
        if cache is not None:
            cache[name] = val
        return val

    inp = hp("all_nodes_hook", tokens)
    dup = hp("hook_duplicate", ioi_duplicate_head(inp))
    sinh = hp("hook_s_inhibition", ioi_s_inhibition_head(inp, dup))
    out = hp("hook_name_mover", ioi_name_mover_head(inp, sinh, names, d_vocab))
    return out


def ioi_hl_intervened(base_tokens, src_tokens, names, d_vocab, hl_node_name):
    """base_model_pair.py:80,89-91 for IOI_HL: source run cached, base run with one hook replaced."""
    cache = {}
    ioi_hl_forward(src_tokens, names, d_vocab, cache=cache)
    return ioi_hl_forward(base_tokens, names, d_vocab, replace={hl_node_name: cache[hl_node_name]})


# --------------------------------------------------------------------------------------------
# LL model 1: torchvision ResNet-18 under HookedModuleWrapper
#   (iit/tasks/mnist_pvr/get_alignment.py:11-15, iit/utils/wrapper.py:6-66)
# State-dict keys follow the wrapper: mod.conv1.mod.weight, mod.layer1.mod.0.mod.bn1.mod.running_mean...
# --------------------------------------------------------------------------------------------
class _HookCtx:
    def __init__(self, hooks, cache, cache_filter):
        self.hooks = hooks or {}
        self.cache = cache
        self.filter = cache_filter

    def __call__(self, name, x):
        if name in self.hooks:
            x = self.hooks[name](x)
        if self.cache is not None and (self.filter is None or name in self.filter):
            self.cache[name] = x.detach()  # [TL-knowledge] cache entries are detached
        return x


class ResNet18Oracle:
    STAGES = [("layer1", 64, 1), ("layer2", 128, 2), ("layer3", 256, 2), ("layer4", 512, 2)]

    def __init__(self, state_dict: dict, momentum=0.1, eps=1e-5):
        self.p = {}
        self.b = {}
        for k, v in state_dict.items():
            v = v.detach().clone().cpu()
            if k.endswith("running_mean") or k.endswith("running_var") or k.endswith("num_batches_tracked"):
                self.b[k] = v
            else:
                self.p[k] = v.float().requires_grad_(True)
        self.training = True
        self.momentum, self.eps = momentum, eps

    def parameters(self):
        return list(self.p.values())

    def state_dict(self):
        d = {k: v.detach().clone() for k, v in self.p.items()}
        d.update({k: v.clone() for k, v in self.b.items()})
        return d

    def train(self, mode=True):
        self.training = mode
        return self

    def eval(self):
        return self.train(False)

    def _bn(self, x, pre):
        if self.training:
            self.b[pre + ".mod.num_batches_tracked"] += 1
        return F.batch_norm(x, self.b[pre + ".mod.running_mean"], self.b[pre + ".mod.running_var"],
                            self.p[pre + ".mod.weight"], self.p[pre + ".mod.bias"],
                            self.training, self.momentum, self.eps)

    def _block(self, x, pre, stride, has_ds, hp):
        identity = x
        out = hp(f"{pre}.mod.conv1.hook_point", F.conv2d(x, self.p[f"{pre}.mod.conv1.mod.weight"], None, stride, 1))
        out = hp(f"{pre}.mod.bn1.hook_point", self._bn(out, f"{pre}.mod.bn1"))
        out = hp(f"{pre}.mod.relu.hook_point", F.relu(out, inplace=True))  # torchvision ReLU(inplace=True)
        out = hp(f"{pre}.mod.conv2.hook_point", F.conv2d(out, self.p[f"{pre}.mod.conv2.mod.weight"], None, 1, 1))
        out = hp(f"{pre}.mod.bn2.hook_point", self._bn(out, f"{pre}.mod.bn2"))
        if has_ds:
            d = f"{pre}.mod.downsample"
            identity = hp(f"{d}.mod.0.hook_point", F.conv2d(x, self.p[f"{d}.mod.0.mod.weight"], None, stride, 0))
            identity = hp(f"{d}.mod.1.hook_point", self._bn(identity, f"{d}.mod.1"))
            identity = hp(f"{d}.hook_point", identity)
        out = out + identity  # torchvision: out += identity
        out = hp(f"{pre}.mod.relu.hook_point", F.relu(out, inplace=True))
        return hp(f"{pre}.hook_point", out)

    def forward(self, x, hooks=None, cache=None, cache_filter=None):
        hp = _HookCtx(hooks, cache, cache_filter)
        x = hp("mod.conv1.hook_point", F.conv2d(x, self.p["mod.conv1.mod.weight"], None, 2, 3))
        x = hp("mod.bn1.hook_point", self._bn(x, "mod.bn1"))
        x = hp("mod.relu.hook_point", F.relu(x, inplace=True))
        x = hp("mod.maxpool.hook_point", F.max_pool2d(x, 3, 2, 1))
        for name, ch, stride in self.STAGES:
            x = self._block(x, f"mod.{name}.mod.0", stride, stride != 1, hp)
            x = self._block(x, f"mod.{name}.mod.1", 1, False, hp)
            x = hp(f"mod.{name}.hook_point", x)
        x = hp("mod.avgpool.hook_point", F.adaptive_avg_pool2d(x, 1))
        x = torch.flatten(x, 1)
        return hp("mod.fc.hook_point", F.linear(x, self.p["mod.fc.mod.weight"], self.p["mod.fc.mod.bias"]))

    __call__ = forward

    @staticmethod
    def conv_hook_points():
        """iit/utils/wrapper.py:69-70 -- hook names containing 'conv' (17)."""
        names = ["mod.conv1.hook_point"]
        for l in range(1, 5):
            for b in range(2):
                for c in (1, 2):
                    names.append(f"mod.layer{l}.mod.{b}.mod.conv{c}.hook_point")
        return names


def pvr_corr(mode: str, hook_point: str, channels: int, dim: int) -> dict:
    """iit/tasks/mnist_pvr/pvr_hl.py:68-134 -- HL hook name -> [(ll hook name, index spec)]."""
    if mode == "c":
        cs = channels // 4
        return {h: [(hook_point, (None, slice(cs * i, cs * (i + 1)) if i else slice(None, cs), None, None))]
                for i, h in enumerate(PVR_HOOKS)}
    q = dim // 2
    boxes = pvr_quadrant_boxes(q)
    return {h: [(hook_point, boxes[i])] for i, h in enumerate(PVR_HOOKS)}


def pvr_quadrant_boxes(q):
    """pvr_hl.py:104-132 / pvr_check_leaky_hl.py:80-89."""
    return [
        (None, None, slice(None, q), slice(None, q)),
        (None, None, slice(None, q), slice(q, q * 2)),
        (None, None, slice(q, q * 2), slice(None, q)),
        (None, None, slice(q, q * 2), slice(q, q * 2)),
    ]


def pvr_leaky_corr(hook_point: str, dim: int) -> dict:
    """pvr_check_leaky_hl.py:19-27,72-109 -- 12 HL nodes, each -> the 'to_' quadrant at hook_point.
    Insertion order follows the reference's nested i/j loops."""
    boxes = pvr_quadrant_boxes(dim // 2)
    corr = {}
    for i in ["tl", "tr", "bl", "br"]:
        for j in ["tl", "tr", "bl", "br"]:
            if i != j:
                name = f"hook_{i}_leaked_to_{j}"
                corr[name] = [(hook_point, boxes[leaky_target_quadrant(name)])]
    return corr


# --------------------------------------------------------------------------------------------
# LL model 2: HookedTransformer (train_ioi.py:30-34, iit/tasks/ioi/__init__.py:8-17) [TL-knowledge]
# --------------------------------------------------------------------------------------------
def gelu_new(x):
    return 0.5 * x * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * torch.pow(x, 3.0))))


class TransformerOracle:
    def __init__(self, state_dict, n_layers, n_heads, d_head, eps=1e-5, act="gelu_new"):
        self.p = {k: v.detach().clone().cpu().float().requires_grad_(True)
                  for k, v in state_dict.items()
                  if not (k.endswith(".mask") or k.endswith(".IGNORE"))}
        self.n_layers, self.n_heads, self.d_head, self.eps = n_layers, n_heads, d_head, eps
        self.act = {"gelu_new": gelu_new, "gelu": F.gelu, "relu": F.relu}[act]
        self.training = True

    def parameters(self):
        return list(self.p.values())

    def train(self, mode=True):
        self.training = mode
        return self

    def eval(self):
        return self.train(False)

    def _lnpre(self, x):
        x = x - x.mean(-1, keepdim=True)
        return x / (x.pow(2).mean(-1, keepdim=True) + self.eps).sqrt()

    def forward(self, tokens, hooks=None, cache=None, cache_filter=None):
        hp = _HookCtx(hooks, cache, cache_filter)
        p = self.p
        b, s = tokens.shape
        resid = p["embed.W_E"][tokens] + p["pos_embed.W_pos"][:s].unsqueeze(0)
        mask = torch.tril(torch.ones(s, s, dtype=torch.bool))
        for l in range(self.n_layers):
            pre = f"blocks.{l}"
            x = self._lnpre(resid)
            q = torch.einsum("bpm,hmd->bphd", x, p[f"{pre}.attn.W_Q"]) + p[f"{pre}.attn.b_Q"]
            k = torch.einsum("bpm,hmd->bphd", x, p[f"{pre}.attn.W_K"]) + p[f"{pre}.attn.b_K"]
            v = torch.einsum("bpm,hmd->bphd", x, p[f"{pre}.attn.W_V"]) + p[f"{pre}.attn.b_V"]
            scores = torch.einsum("bqhd,bkhd->bhqk", q, k) / math.sqrt(self.d_head)
            scores = torch.where(mask, scores, torch.tensor(-torch.inf))
            pattern = F.softmax(scores, dim=-1)
            z = hp(f"{pre}.attn.hook_z", torch.einsum("bkhd,bhqk->bqhd", v, pattern))
            attn_out = torch.einsum("bqhd,hdm->bqm", z, p[f"{pre}.attn.W_O"]) + p[f"{pre}.attn.b_O"]
            resid = resid + attn_out
            x = self._lnpre(resid)
            h = x @ p[f"{pre}.mlp.W_in"] + p[f"{pre}.mlp.b_in"]
            post = hp(f"{pre}.mlp.hook_post", self.act(h))
            resid = resid + post @ p[f"{pre}.mlp.W_out"] + p[f"{pre}.mlp.b_out"]
        x = self._lnpre(resid)
        return x @ p["unembed.W_U"] + p["unembed.b_U"]

    __call__ = forward


IOI_CORR = {  # iit/tasks/ioi/__init__.py:18-27, in dict order
    "hook_duplicate": ["blocks.0.attn.hook_z"],
    "hook_s_inhibition": ["blocks.2.attn.hook_z", "blocks.3.attn.hook_z"],
    "hook_name_mover": ["blocks.4.attn.hook_z", "blocks.5.attn.hook_z"],
    "all_nodes_hook": ["blocks.0.mlp.hook_post", "blocks.1.mlp.hook_post"],
}


def ioi_corr() -> dict:
    """correspondence.py:60-70 make_corr_from_dict: every LL node has index=None (whole tensor)."""
    return {k: [(n, None) for n in v] for k, v in IOI_CORR.items()}


def all_nodes(n_layers, n_heads, suffixes):
    """iit/utils/node_picker.py:9-26 -- per layer: one node per head [:, :, h, :] then the MLP whole."""
    nodes = []
    for layer in range(n_layers):
        for head in range(n_heads):
            nodes.append((f"blocks.{layer}.{suffixes['attn']}", (slice(None), slice(None), head, slice(None))))
        nodes.append((f"blocks.{layer}.{suffixes['mlp']}", None))
    return nodes


def nodes_not_in_circuit(n_layers, n_heads, corr, suffixes):
    """node_picker.py:43-54 with index.py:80-112: a node is dropped when it shares a hook name with an
    in-circuit node whose index intersects (whole-tensor index intersects everything)."""
    in_circuit = [n for nodes in corr.values() for n in nodes]
    out = []
    for name, spec in all_nodes(n_layers, n_heads, suffixes):
        hit = False
        for cname, cspec in in_circuit:
            if cname == name and specs_intersect(spec, cspec):
                hit = True
        if not hit:
            out.append((name, spec))
    return out


def specs_intersect(a, b) -> bool:
    """iit/utils/index.py:80-112."""
    if a is None or b is None:
        return True
    ai, bi = as_index(a), as_index(b)
    if len(ai) != len(bi):
        raise ValueError("Cannot compare indices of different lengths")
    for x, y in zip(ai, bi):
        if x == slice(None) or y == slice(None):
            continue
        if isinstance(x, int) and isinstance(y, int):
            if x != y:
                return False
        elif isinstance(x, slice) and isinstance(y, slice):
            start = 0 if (x.start is None and y.start is None) else max(x.start or 0, y.start or 0)
            stop = min(x.stop if x.stop is not None else math.inf, y.stop if y.stop is not None else math.inf)
            if start >= stop:
                return False
        else:
            sl, i = (y, x) if isinstance(x, int) else (x, y)
            lo = sl.start if sl.start is not None else 0
            hi = sl.stop if sl.stop is not None else math.inf
            if not (lo <= i < hi):
                return False
    return True


# --------------------------------------------------------------------------------------------
# L3: intervention, losses, steps
# --------------------------------------------------------------------------------------------
def ll_intervention(ll, base_x, src_x, ll_nodes):
    """base_model_pair.py:86-98 -- source run cached (detached), base run with one patch per LL node."""
    cache = {}
    ll(src_x, cache=cache, cache_filter={n for n, _ in ll_nodes})
    hooks = {}
    for name, spec in ll_nodes:
        hooks[name] = (lambda act, name=name, spec=spec: patch(act, cache[name], spec))
    return ll(base_x, hooks=hooks), cache


def iia_counts(ll_output, hl_output):
    """iit_behavior_model_pair.py:118-120 -- (#argmax(ll)==hl, batch). argmax = first maximal index."""
    top1 = torch.argmax(ll_output, dim=-1)
    return int((top1 == hl_output).sum()), int(hl_output.numel())


def causality_stat(ll_output, hl_output, base_y, src_y):
    """eval_causality.py:74-83 -- sum(acc*changed) / (sum(changed)+1e-10) for one (batch, node)."""
    changed = (src_y != base_y).float()
    top1 = torch.argmax(ll_output, dim=1)
    acc = (top1 == hl_output).float() * changed
    return acc.sum() / (changed.sum() + 1e-10)


class OraclePair:
    """Restates IITBehaviorModelPair / StrictIITModelPair / IOI_ModelPair step logic
    (iit_behavior_model_pair.py:51-141, strict_iit_model_pair.py:36-91, ioi_model_pair.py:22-112,
    base_model_pair.py:117-118,176-180) over an oracle LL model. `task` is 'pvr' or 'ioi'."""

    def __init__(self, ll, task, corr, *, seed=0, lr=1e-3, clip=1.0, weights=(1.0, 1.0, 1.0),
                 strict_nodes=None, names=None, d_vocab=None, use_single_loss=False):
        self.ll, self.task, self.corr = ll, task, corr
        self.rng = np.random.default_rng(seed)  # iit_model_pair.py:29
        self.opt = torch.optim.Adam(ll.parameters(), lr=lr)  # base_model_pair.py:229
        self.clip = clip
        self.w_iit, self.w_beh, self.w_strict = weights
        self.strict_nodes = strict_nodes
        self.names, self.d_vocab = names, d_vocab
        self.use_single_loss = use_single_loss
        self.keys = list(corr.keys())

    # base_model_pair.py:117-118 / strict_iit_model_pair.py:33-34
    def sample_hl_name(self):
        return str(self.rng.choice(self.keys))

    def sample_ll_node(self):
        return self.strict_nodes[int(self.rng.choice(len(self.strict_nodes)))]

    def hl_out(self, base, src, node):
        if self.task == "pvr":
            return pvr_hl_intervened(base[2], src[2], node)
        return ioi_hl_intervened(base[0], src[0], self.names, self.d_vocab, node)

    def loss_fn(self, out, target):
        """ioi_model_pair.py:27-35 (next_token is always False) / iit_behavior:20-30 CrossEntropyLoss."""
        if out.dim() == 3:
            return F.cross_entropy(out[:, -1, :], target[:, -1, :])
        return F.cross_entropy(out, target)

    def do_intervention(self, base, src, node):
        ll_out, _ = ll_intervention(self.ll, base[0], src[0], self.corr[node])
        return self.hl_out(base, src, node), ll_out

    def iit_loss(self, base, src, node):
        hl, ll_out = self.do_intervention(base, src, node)
        if self.task == "ioi":  # ioi_model_pair.py:55-69
            am = torch.argmax(hl[:, -1, :], dim=-1)
            probs = F.one_hot(am, num_classes=hl.shape[-1]).float()
            return self.loss_fn(ll_out[:, -1, :], probs)
        return self.loss_fn(ll_out, hl)  # base_model_pair.py:165-174

    def step_on_loss(self, loss):  # iit_behavior_model_pair.py:60-64
        self.opt.zero_grad()
        loss.backward()
        if self.clip:
            torch.nn.utils.clip_grad_norm_(self.ll.parameters(), self.clip)
        self.opt.step()

    def train_step(self, base, src, strict=False):
        self.ll.train()
        out = {}
        node = self.sample_hl_name()
        iit = self.iit_loss(base, src, node) * self.w_iit
        if not self.use_single_loss:
            self.step_on_loss(iit)
        total = iit
        if strict:  # strict_iit_model_pair.py:57-74
            ll_node = self.sample_ll_node()
            o, _ = ll_intervention(self.ll, base[0], src[0], [ll_node])
            sl = self.loss_fn(o.squeeze(), base[1]) * self.w_strict
            if not self.use_single_loss:
                self.step_on_loss(sl)
            total = total + sl
            out["train/strict_loss"] = sl.item()
            out["ll_node"] = ll_node
        o = self.ll(base[0])  # iit_behavior_model_pair.py:51-58
        beh = self.loss_fn(o.squeeze(), base[1]) * self.w_beh
        if not self.use_single_loss:
            self.step_on_loss(beh)
        else:
            self.step_on_loss(total + beh)
        out.update({"train/iit_loss": iit.item(), "train/behavior_loss": beh.item(), "hl_node": node})
        return out

    @torch.no_grad()
    def eval_step(self, base, src):
        self.ll.eval()
        node = self.sample_hl_name()
        hl, ll_out = self.do_intervention(base, src, node)
        if self.task == "ioi":  # ioi_model_pair.py:71-112
            am = torch.argmax(hl[:, -1, :], dim=-1)
            loss = self.loss_fn(ll_out[:, -1, :], F.one_hot(am, num_classes=hl.shape[-1]).float())
            top1 = torch.argmax(ll_out, dim=-1)
            hits = int((top1[:, -1] == am).sum())
            iia = (top1[:, -1] == torch.argmax(hl, dim=-1)[:, -1]).float().mean().item()
            o = self.ll(base[0])
            y = torch.argmax(base[1], dim=-1) if base[1].shape == o.shape else base[1]
            pta = (torch.argmax(o, dim=-1) == y).float().mean(dim=0).numpy()
            return {"val/iit_loss": loss.item(), "val/IIA": iia, "val/accuracy": pta[-1],
                    "val/per_token_accuracy": pta, "hl_node": node, "iia_hits": hits}
        loss = self.loss_fn(ll_out, hl)  # iit_behavior_model_pair.py:102-141
        hits, n = iia_counts(ll_out, hl)
        iia = (torch.argmax(ll_out, dim=-1) == hl).float().mean().item()
        o = self.ll(base[0])
        acc = (torch.argmax(o, dim=-1) == base[1]).float().mean().item()
        return {"val/iit_loss": loss.item(), "val/IIA": iia, "val/accuracy": acc, "hl_node": node,
                "iia_hits": hits}


# --------------------------------------------------------------------------------------------
# Freeze / stop-gradient pairs AS WRITTEN in the reference
# --------------------------------------------------------------------------------------------
class _StopGradSite(torch.autograd.Function):
    """One hook point under StopGradHookedModel (stop_grad_pair.py:37-44,63-77,79-103): forward `act / scale`; backward:
    the incoming gradient with every listed index zeroed, handed on WITHOUT the 1/scale factor -- transformer_lens wires a
    backward hook as a full backward hook of the HookPoint module, whose return value (`[grad]`, the hook's edited copy of
    the gradient w.r.t. the module OUTPUT) replaces the gradient w.r.t. the module INPUT, skipping the division. [TL-knowledge]"""

    @staticmethod
    def forward(ctx, act, scale, specs):
        ctx.specs = specs
        return act / scale if scale else act.clone()

    @staticmethod
    def backward(ctx, g):
        g = g.clone()
        for spec in ctx.specs:
            g[as_index(spec)] = 0
        return g, None, None


class OracleFreezePair:
    """FreezedModelPair (stop_grad=False) and StopGradModelPair (stop_grad=True) over a TransformerOracle, restating what
    the reference code does as written:
      * both are IITBehaviorModelPair steps (iit_behavior_model_pair.py:66-141): IIT term then behaviour term, one
        clip + Adam update each; the loss is CrossEntropyLoss on the full [B,S,V] logits, i.e. over torch's class
        dimension 1, against the dense HL logits / the one-hot labels as probability targets;
      * freeze_model_pair.py:23-38: the "zeroing" assigns .grad on a temporary and changes nothing;
      * stop_grad_pair.py:79-103: `ll_model(x)` (the behaviour forward, the accuracy forward of eval) attaches the
        scale / zero-gradient hooks and LEAVES them attached; they act once more, on the next run_with_cache (the source
        run of the next intervention), which removes them when it ends."""

    def __init__(self, ll, corr, names, d_vocab, stop_grad, n_layers=6, n_heads=4, scale=1e6, seed=0, lr=1e-3, clip=1.0,
                 use_ln_hooks=True):
        self.ll, self.corr, self.names, self.d_vocab = ll, corr, names, d_vocab
        self.rng = np.random.default_rng(seed)
        self.opt = torch.optim.Adam(ll.parameters(), lr=lr)
        self.clip, self.stop_grad, self.scale, self.use_ln_hooks = clip, stop_grad, scale, use_ln_hooks
        self.keys = list(corr.keys())
        outside = nodes_not_in_circuit(n_layers, n_heads, corr, {"attn": "attn.hook_z", "mlp": "mlp.hook_post"})
        self.sites = {}                       # hook name -> index specs zeroed there (node_picker.py:57-79 post nodes)
        for name, spec in outside:
            self.sites.setdefault(name, []).append(spec)
        self.leftover = False

    def _site_hooks(self):
        sc = self.scale if self.use_ln_hooks else None
        return {name: (lambda act, specs=tuple(specs): _StopGradSite.apply(act, sc, specs)) for name, specs in self.sites.items()}

    def model_call(self, x):
        """`self.ll_model(x)`: StopGradHookedModel.forward, or the plain model for the freeze pair."""
        if not self.stop_grad:
            return self.ll(x)
        self.leftover = True
        return self.ll(x, hooks=self._site_hooks())

    def do_intervention(self, base, src, node):
        hooks = self._site_hooks() if self.leftover else None
        self.leftover = False                 # run_with_cache's exit clears the level the leftovers live on
        cache = {}
        ll_nodes = self.corr[node]
        self.ll(src[0], hooks=hooks, cache=cache, cache_filter={n for n, _ in ll_nodes})
        patches = {n: (lambda act, n=n, sp=sp: patch(act, cache[n], sp)) for n, sp in ll_nodes}
        return ioi_hl_intervened(base[0], src[0], self.names, self.d_vocab, node), self.ll(base[0], hooks=patches)

    def step_on_loss(self, loss):
        self.opt.zero_grad()
        loss.backward()
        if self.clip:
            torch.nn.utils.clip_grad_norm_(self.ll.parameters(), self.clip)
        self.opt.step()

    def iit_loss(self, base, src, node):
        hl, ll_out = self.do_intervention(base, src, node)
        return F.cross_entropy(ll_out, hl)

    def behaviour_loss(self, base):
        return F.cross_entropy(self.model_call(base[0]).squeeze(), base[1])

    def train_step(self, base, src):
        self.ll.train()
        node = str(self.rng.choice(self.keys))
        iit = self.iit_loss(base, src, node)
        self.step_on_loss(iit)
        beh = self.behaviour_loss(base)
        self.step_on_loss(beh)
        return {"train/iit_loss": iit.item(), "train/behavior_loss": beh.item(), "hl_node": node}

    @torch.no_grad()
    def eval_step(self, base, src):
        self.ll.eval()
        node = str(self.rng.choice(self.keys))
        hl, ll_out = self.do_intervention(base, src, node)
        loss = F.cross_entropy(ll_out, hl)
        iia = (torch.argmax(ll_out, dim=-1) == torch.argmax(hl, dim=-1)).float().mean().item()
        out = self.model_call(base[0])
        acc = (torch.argmax(out, dim=-1) == torch.argmax(base[1], dim=-1)).float().mean().item()
        return {"val/iit_loss": loss.item(), "val/IIA": iia, "val/accuracy": acc, "hl_node": node}


# --------------------------------------------------------------------------------------------
# Synthetic inputs; shared by tests, smoke and bench so both arms see the
# same pairs. Pure torch-CPU generators: deterministic for a given torch build.
# --------------------------------------------------------------------------------------------
def synthetic_pvr(batch, seed, size=56):
    g = torch.Generator().manual_seed(seed)
    tg = torch.Generator().manual_seed(1234)
    q = size // 2
    templates = torch.rand(10, q, q, generator=tg)
    vars_ = torch.randint(0, 10, (batch, 4), generator=g)
    noise = torch.rand(batch, 4, q, q, generator=g) * 0.25
    quads = (templates[vars_] * 0.75 + noise)  # [B,4,q,q] in [0,1]
    top = torch.cat([quads[:, 0], quads[:, 1]], dim=-1)
    bot = torch.cat([quads[:, 2], quads[:, 3]], dim=-1)
    img = torch.cat([top, bot], dim=-2).unsqueeze(1).repeat(1, 3, 1, 1).contiguous()
    cm = torch.tensor(MNIST_CLASS_MAP)
    y = vars_[torch.arange(batch), cm[vars_[:, 0]]]  # dataset.py:109-111
    return img, y, vars_


def synthetic_pvr_counterfactual(base, hl_node_name, seed):
    """dataset.py:125-176 restated on tensors: the source differs from the base in exactly the
    quadrant of the hooked variable, with a different class."""
    x, y, v = base
    b, _, size, _ = x.shape
    q = size // 2
    i = pvr_var_index(hl_node_name)
    g = torch.Generator().manual_seed(seed)
    tg = torch.Generator().manual_seed(1234)
    templates = torch.rand(10, q, q, generator=tg)
    new_cls = (v[:, i] + torch.randint(1, 10, (b,), generator=g)) % 10
    noise = torch.rand(b, q, q, generator=g) * 0.25
    quad = templates[new_cls] * 0.75 + noise
    r0, c0 = (i // 2) * q, (i % 2) * q
    nx = x.clone()
    nx[:, :, r0:r0 + q, c0:c0 + q] = quad.unsqueeze(1)
    nv = v.clone()
    nv[:, i] = new_cls
    cm = torch.tensor(MNIST_CLASS_MAP)
    ny = nv[torch.arange(b), cm[nv[:, 0]]]
    return nx, ny, nv


def synthetic_ioi(batch, seed, d_vocab=50257, seq=16, n_names=100, one_hot=True):
    """Template `BOS T , A and B went to the L . S gave the O to` -> 17 ids; x = ids[:-1] (16),
    y = ids[1:]; IO = the name that is not S (ioi_dataset_tl.py:372-377)."""
    g = torch.Generator().manual_seed(seed)
    ng = torch.Generator().manual_seed(4321)
    lo = min(1000, d_vocab // 4)
    names = (torch.randperm(max(d_vocab - lo - 1, n_names), generator=ng)[:n_names] + lo).sort().values
    n_names = len(names)
    fixed = [(7 + 13 * i) % lo for i in range(seq + 1)]
    a = torch.randint(0, n_names, (batch,), generator=g)
    bidx = (a + torch.randint(1, n_names, (batch,), generator=g)) % n_names
    s_is_a = torch.randint(0, 2, (batch,), generator=g).bool()
    ids = torch.tensor(fixed).repeat(batch, 1)
    ids[:, 0] = d_vocab - 1
    A, Bn = names[a], names[bidx]
    S = torch.where(s_is_a, A, Bn)
    IO = torch.where(s_is_a, Bn, A)
    ids[:, 3], ids[:, 5], ids[:, 11], ids[:, seq] = A, Bn, S, IO
    x = ids[:, :seq].contiguous()
    y_ids = ids[:, 1:seq + 1].contiguous()
    y = F.one_hot(y_ids, num_classes=d_vocab).float() if one_hot else y_ids
    return (x, y, IO.unsqueeze(1)), names


def resnet18_state_dict(seed=0):
    """Seeded torchvision-equivalent ResNet-18 init (fc -> 10) in the wrapper's key names
    (get_alignment.py:11-15): conv kaiming_normal(fan_out, relu); BN weight 1 / bias 0;
    fc: nn.Linear default init. Drawn in module-definition order from one generator."""
    g = torch.Generator().manual_seed(seed)
    sd = {}

    def conv(key, cout, cin, k):
        std = math.sqrt(2.0 / (cout * k * k))
        sd[key + ".mod.weight"] = torch.randn(cout, cin, k, k, generator=g) * std

    def bn(key, c):
        sd[key + ".mod.weight"] = torch.ones(c)
        sd[key + ".mod.bias"] = torch.zeros(c)
        sd[key + ".mod.running_mean"] = torch.zeros(c)
        sd[key + ".mod.running_var"] = torch.ones(c)
        sd[key + ".mod.num_batches_tracked"] = torch.tensor(0, dtype=torch.long)

    conv("mod.conv1", 64, 3, 7)
    bn("mod.bn1", 64)
    cin = 64
    for name, ch, stride in ResNet18Oracle.STAGES:
        for blk in range(2):
            pre = f"mod.{name}.mod.{blk}"
            s = stride if blk == 0 else 1
            conv(pre + ".mod.conv1", ch, cin, 3)
            bn(pre + ".mod.bn1", ch)
            conv(pre + ".mod.conv2", ch, ch, 3)
            bn(pre + ".mod.bn2", ch)
            if blk == 0 and s != 1:
                conv(pre + ".mod.downsample.mod.0", ch, cin, 1)
                bn(pre + ".mod.downsample.mod.1", ch)
            cin = ch
    bound = 1.0 / math.sqrt(512)
    sd["mod.fc.mod.weight"] = (torch.rand(10, 512, generator=g) * 2 - 1) * bound
    sd["mod.fc.mod.bias"] = (torch.rand(10, generator=g) * 2 - 1) * bound
    return sd


def transformer_state_dict(seed=0, n_layers=6, n_heads=4, d_model=64, d_head=16, d_mlp=3072,
                           d_vocab=50257, n_ctx=1024, std=0.1):
    """normal(0, std) W's, zero biases; TL parameter names [TL-knowledge]."""
    g = torch.Generator().manual_seed(seed)
    sd = {}

    def w(key, *shape):
        sd[key] = torch.randn(*shape, generator=g) * std

    w("embed.W_E", d_vocab, d_model)
    w("pos_embed.W_pos", n_ctx, d_model)
    for l in range(n_layers):
        p = f"blocks.{l}"
        for n in ("W_Q", "W_K", "W_V"):
            w(f"{p}.attn.{n}", n_heads, d_model, d_head)
        w(f"{p}.attn.W_O", n_heads, d_head, d_model)
        for n in ("b_Q", "b_K", "b_V"):
            sd[f"{p}.attn.{n}"] = torch.zeros(n_heads, d_head)
        sd[f"{p}.attn.b_O"] = torch.zeros(d_model)
        w(f"{p}.mlp.W_in", d_model, d_mlp)
        sd[f"{p}.mlp.b_in"] = torch.zeros(d_mlp)
        w(f"{p}.mlp.W_out", d_mlp, d_model)
        sd[f"{p}.mlp.b_out"] = torch.zeros(d_model)
    w("unembed.W_U", d_model, d_vocab)
    sd["unembed.b_U"] = torch.zeros(d_vocab)
    return sd


def ioi_ablation_fixture(bias: float, n=48, d_vocab=101, n_names=4):
    """Weights, samples and names of the ablation-evaluation fixtures (oracle/make_golden.py:golden_ioi_ablations):
    the small IOI transformer with `bias` added to one name's unembed bias, `n` prompts over `n_names` names."""
    sd = transformer_state_dict(0, d_mlp=128, d_vocab=d_vocab, n_ctx=32)
    (x, y, iv), names = synthetic_ioi(n, 100, d_vocab=d_vocab, n_names=n_names)
    sd = dict(sd)
    b = sd["unembed.b_U"].clone()
    b[names[1]] += bias
    sd["unembed.b_U"] = b
    data = [(x[i], y[i], iv[i]) for i in range(n)]
    return sd, data, names


# --------------------------------------------------------------------------------------------
# ablation evaluations (iit/utils/eval_ablations.py, eval_metrics.py) -- SURVEY.md 8(f) row 2
# --------------------------------------------------------------------------------------------
def kl_div_rows(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """eval_metrics.py:4-20 -- rows that do not sum to 1 are softmaxed; KL(b || a) summed over classes."""
    def pmf(x):
        s = x.sum(dim=-1)
        return x if torch.allclose(s, torch.ones_like(s)) else F.softmax(x, dim=-1)
    return F.kl_div(pmf(a).log(), pmf(b), reduction="none", log_target=False).sum(dim=-1)


def pair_batches(n_items, dataset_seed, batch_size, n_base=None, n_abl=None):
    """Batches of (base row, ablation row) exactly as IITDataset + DataLoader(shuffle=True) form them
    (iit_dataset.py:24-72): per-item numpy pair rule, item order from torch's global RNG (seed it before calling)."""
    n_base = n_items if n_base is None else n_base
    n_abl = n_items if n_abl is None else n_abl
    table = []
    for i in range(n_items):
        rng = np.random.default_rng(dataset_seed * 1000000 + i)
        table.append((int(rng.choice(n_base)), int(rng.choice(n_abl))))
    loader = torch.utils.data.DataLoader(range(n_items), batch_size=batch_size, shuffle=True, collate_fn=list)
    return [[table[i] for i in items] for items in loader]


def unique_batches(n_items, batch_size):
    """Item batches of IITUniqueDataset.make_loader (eval_datasets.py:7-19): the same shuffled DataLoader over rows."""
    return list(torch.utils.data.DataLoader(range(n_items), batch_size=batch_size, shuffle=True, collate_fn=list))


def resample_kl_effects(ll, names, d_vocab, data, nodes, batch_size, dataset_seed=0):
    """check_causal_effect with the KL metric (eval_ablations.py:35-61,128-161) for IOI: label index = last position.
    `nodes`: [(hook name, index spec)]. Returns one float per node (mean over batches of the batch-mean KL)."""
    x = torch.stack([d[0] for d in data])
    batches = pair_batches(len(data), dataset_seed, batch_size)
    out = [0.0] * len(nodes)
    with torch.no_grad():
        for rows in batches:
            bx = x[[r[0] for r in rows]]
            ax = x[[r[1] for r in rows]]
            hl = ioi_hl_forward(bx, names, d_vocab)[:, -1, :]
            for k, node in enumerate(nodes):
                ll_out, _ = ll_intervention(ll, bx, ax, [node])
                out[k] += kl_div_rows(ll_out[:, -1, :], hl).mean().item()
    return [v / len(batches) for v in out]


def mean_cache(ll, data, names, batch_size):
    """get_mean_cache (eval_ablations.py:164-173): sum over batches of batch-mean / number of batches."""
    x = torch.stack([d[0] for d in data])
    batches = unique_batches(len(data), batch_size)
    out = {}
    with torch.no_grad():
        for items in batches:
            cache = {}
            ll(x[items], cache=cache, cache_filter=set(names))
            for name, tensor in cache.items():
                if name not in out:
                    out[name] = torch.zeros_like(tensor[0].unsqueeze(0))
                out[name] += tensor.mean(dim=0).unsqueeze(0) / len(batches)
    return out


def ablated_logits(ll, x, node, mean=None):
    """Base run with the node's slice overwritten by the dataset mean, or zeros (eval_ablations.py:176-193)."""
    name, spec = node
    idx = as_index(spec)

    def hook(act):
        out = act.clone()
        out[idx] = 0 if mean is None else mean[name].expand_as(act)[idx]
        return out

    with torch.no_grad():
        return ll(x, hooks={name: hook})


def ablation_effects(ll, names, d_vocab, data, nodes, batch_size, mean=None):
    """check_causal_effect_on_ablation (eval_ablations.py:196-237,262-295): per node, mean over batches of
    #(correct before, prediction != HL after) / (#correct before + 1e-6)."""
    x = torch.stack([d[0] for d in data])
    batches = unique_batches(len(data), batch_size)
    out = [0.0] * len(nodes)
    with torch.no_grad():
        for items in batches:
            bx = x[items]
            hl_top = ioi_hl_forward(bx, names, d_vocab)[:, -1, :].argmax(-1)
            base_top = ll(bx)[:, -1, :].argmax(-1)
            acc = (base_top == hl_top).float()
            for k, node in enumerate(nodes):
                top = ablated_logits(ll, bx, node, mean)[:, -1, :].argmax(-1)
                out[k] += (((top != hl_top).float() * acc).sum().item()) / (acc.sum().item() + 1e-6)
    return [v / len(batches) for v in out]
