"""Low-level model for the IOI task: a small GPT-2-style `HookedTransformer` whose layers are the
sm_100a kernels.

Replaces `transformer_lens.HookedTransformer(cfg)` as `train_ioi.py:30-34` builds it (gpt2 config with
iit/tasks/ioi/__init__.py:8-17 applied: 6 layers, 4 heads, d_model 64, d_head 16; inherited d_mlp 3072,
d_vocab 50257, n_ctx 1024, gelu_new, LNPre = folded layer norm) and keeps what callers see [TL-knowledge]:
  * parameter names / shapes: embed.W_E [V,D], pos_embed.W_pos [n_ctx,D], blocks.i.attn.W_Q/K/V [H,D,dh],
    W_O [H,dh,D], b_Q/K/V [H,dh], b_O [D], blocks.i.mlp.W_in [D,d_mlp], b_in, W_out [d_mlp,D], b_out,
    unembed.W_U [D,V], b_U (+ the attention `mask` / `IGNORE` buffers, so reference checkpoints load);
  * hook-point names: blocks.i.attn.hook_z [B,S,H,dh] and blocks.i.mlp.hook_post [B,S,d_mlp] (the patch
    sites of the shipped correspondence and of the strict loss), plus hook_resid_pre/mid/post,
    hook_attn_out, hook_mlp_out, mlp.hook_pre. The remaining transformer_lens hook points exist by name;
    putting a hook on one of those raises (their tensors live only inside fused kernels);
  * `cfg` with n_layers, n_heads, d_vocab_out, device, to_dict().
`forward(tokens, last_pos_only=True)` unembeds only the last position ([B,V]): all the IOI losses read
nothing else (ioi_model_pair.py:20,35), and it avoids the [B,16,50257] logits tensor.
"""
from __future__ import annotations

import math
from types import SimpleNamespace

import torch
import torch.nn as nn

from .. import ops
from ..hooks import HookedRootModule, HookPoint


class TransformerConfig(SimpleNamespace):
    def to_dict(self):
        return dict(self.__dict__)


def make_cfg(cfg) -> TransformerConfig:
    d = dict(n_layers=None, d_model=None, n_ctx=1024, d_head=None, n_heads=-1, d_mlp=None, act_fn="gelu_new",
             d_vocab=50257, d_vocab_out=-1, eps=1e-5, normalization_type="LNPre", init_weights=True, initializer_range=-1.0,
             seed=None, device=None, use_attn_result=False, attn_only=False, dtype=torch.float32)
    d.update(dict(cfg) if isinstance(cfg, dict) else cfg.__dict__)
    c = TransformerConfig(**d)
    if c.n_heads == -1:
        c.n_heads = c.d_model // c.d_head
    if c.d_vocab_out == -1:
        c.d_vocab_out = c.d_vocab
    if c.d_mlp is None:
        c.d_mlp = 4 * c.d_model
    if c.initializer_range < 0:
        c.initializer_range = 0.8 / math.sqrt(c.d_model)
    if c.device is None:
        c.device = "cuda" if torch.cuda.is_available() else "cpu"
    if c.normalization_type != "LNPre" or c.act_fn != "gelu_new" or c.use_attn_result or c.attn_only:
        raise NotImplementedError("the kernels implement the train_ioi.py configuration: LNPre, gelu_new, "
                                  "use_attn_result=False, attn_only=False")
    return c


class _Holder(nn.Module):
    """Namespace module so that parameters / hook points get transformer_lens' dotted names."""


DIRECT_WEIGHT_GRADS = True    # weight-gradient GEMMs add into the optimizer's gradient views (tests flip it to compare)
FUSE_GELU_IN_GEMM = False   # True: gelu_new / its backward in the GEMM epilogues (iit_gemm_bias_act_tc act=1, iit_gemm_dgrad_gelu_tc)


def _hp():
    # every hook point this model lets a hook modify (hook_z, mlp.hook_post, the residual-stream points, hook_attn_out /
    # hook_mlp_out) carries a tensor whose producer kept its INPUTS for the backward (attention keeps qkv, bias+GELU keeps
    # the pre-activation, bias+residual keeps nothing), so a patch may overwrite it in place
    return HookPoint(inplace_ok=True)


class HookedTransformerLL(HookedRootModule):
    supports_last_pos = True
    forward_is_pure = True      # no statistics, no dropout: abandoning a forward early changes nothing (base_model_pair.ll_source_run)

    def __init__(self, cfg, dtype=None):
        super().__init__()
        self.cfg = c = make_cfg(cfg)
        self.act_dtype = dtype or torch.float32
        V, D, H, dh, F_ = c.d_vocab, c.d_model, c.n_heads, c.d_head, c.d_mlp
        self.embed = _Holder()
        self.embed.W_E = nn.Parameter(torch.empty(V, D))
        self.hook_embed = _hp()
        self.pos_embed = _Holder()
        self.pos_embed.W_pos = nn.Parameter(torch.empty(c.n_ctx, D))
        self.hook_pos_embed = _hp()
        self.blocks = nn.ModuleList()
        for _ in range(c.n_layers):
            b = _Holder()
            b.ln1 = _Holder(); b.ln1.hook_scale = _hp(); b.ln1.hook_normalized = _hp()
            b.ln2 = _Holder(); b.ln2.hook_scale = _hp(); b.ln2.hook_normalized = _hp()
            a = b.attn = _Holder()
            a.W_Q = nn.Parameter(torch.empty(H, D, dh)); a.W_K = nn.Parameter(torch.empty(H, D, dh))
            a.W_V = nn.Parameter(torch.empty(H, D, dh)); a.W_O = nn.Parameter(torch.empty(H, dh, D))
            a.b_Q = nn.Parameter(torch.zeros(H, dh)); a.b_K = nn.Parameter(torch.zeros(H, dh))
            a.b_V = nn.Parameter(torch.zeros(H, dh)); a.b_O = nn.Parameter(torch.zeros(D))
            a.register_buffer("mask", torch.tril(torch.ones(c.n_ctx, c.n_ctx, dtype=torch.bool)))
            a.register_buffer("IGNORE", torch.tensor(-torch.inf))
            for n in ("hook_k", "hook_q", "hook_v", "hook_z", "hook_attn_scores", "hook_pattern", "hook_result"):
                setattr(a, n, _hp())
            m = b.mlp = _Holder()
            m.W_in = nn.Parameter(torch.empty(D, F_)); m.b_in = nn.Parameter(torch.zeros(F_))
            m.W_out = nn.Parameter(torch.empty(F_, D)); m.b_out = nn.Parameter(torch.zeros(D))
            m.hook_pre = _hp(); m.hook_post = _hp()
            for n in ("hook_attn_in", "hook_q_input", "hook_k_input", "hook_v_input", "hook_mlp_in", "hook_attn_out",
                      "hook_mlp_out", "hook_resid_pre", "hook_resid_mid", "hook_resid_post"):
                setattr(b, n, _hp())
            self.blocks.append(b)
        self.ln_final = _Holder(); self.ln_final.hook_scale = _hp(); self.ln_final.hook_normalized = _hp()
        self.unembed = _Holder()
        self.unembed.W_U = nn.Parameter(torch.empty(D, c.d_vocab_out))
        self.unembed.b_U = nn.Parameter(torch.zeros(c.d_vocab_out))
        if c.init_weights:
            if c.seed is not None:
                torch.manual_seed(c.seed)
            for name, p in self.named_parameters():
                if "W_" in name:
                    nn.init.normal_(p, std=c.initializer_range)
        self.setup()
        # data-parallel gradient buckets (iit_b200/dist.py): the gradient w.r.t. the residual stream entering the middle
        # block exists once the later blocks and the unembedding have theirs
        if c.n_layers >= 2:
            self.dp_bucket_tap = (f"blocks.{c.n_layers // 2}.hook_resid_pre", f"blocks.{c.n_layers // 2}.attn.W_Q")
        # attn.hook_result is not in this list: with use_attn_result=False (the only configuration this model accepts)
        # transformer_lens never calls that hook point, so a hook attached there is legal and never fires -- which is
        # what the reference's `check_causal_effect(node_type="a")` relies on (iit/utils/node_picker.py:11-14 default
        # suffix); the same holds here.
        self._unsupported = [n for n in self.hook_dict if n.split(".")[-1] in (
            "hook_embed", "hook_pos_embed", "hook_scale", "hook_normalized", "hook_k", "hook_q", "hook_v",
            "hook_attn_scores", "hook_pattern", "hook_attn_in", "hook_q_input", "hook_k_input",
            "hook_v_input", "hook_mlp_in")]

    def load_state_dict(self, state_dict, strict=True, assign=False):
        sd = dict(state_dict)
        own = self.state_dict()
        for k in own:  # mask / IGNORE buffers are constants: tolerate checkpoints without them
            if k not in sd and (k.endswith(".mask") or k.endswith(".IGNORE")):
                sd[k] = own[k]
        return super().load_state_dict(sd, strict=strict, assign=assign)

    # --- bf16 GEMM operands of every linear map, refreshed by ONE kernel launch ----------------------------------
    # Each GEMM reads a bf16 [out][in] copy of its weight in the forward and a [in][out] copy in dgrad. Making them per
    # GEMM and pass costs ~200 re-layout launches per train step (11 % of it); here they live in persistent buffers
    # described by a device-resident iit_relayout_t table that one iit_filter_relayout_batched launch rewrites after
    # every optimizer update (FlatAdam.post_step, also inside captured graphs) -- the ResNet's scheme. The fused QKV
    # matrix is assembled by six entries straight from the per-head parameters; the unembedding keeps its zero padding
    # to a multiple of 64 columns because the entries never write it.
    def _linear_sites(self):
        c = self.cfg
        D, H, dh, F_, V = c.d_model, c.n_heads, c.d_head, c.d_mlp, c.d_vocab_out
        sites = []
        for i, blk in enumerate(self.blocks):
            a, m = blk.attn, blk.mlp
            hd = H * dh
            # QKV: W_p[h, d, e] -> oi[(p*hd + h*dh + e)][d] (KRSC with K=H, R=dh, C=D) and io[d][p*hd + h*dh + e] (CRSK
            # with C=D, R=H, S=dh, K=1, rows 3*hd apart)
            ent = []
            for j, w in enumerate((a.W_Q, a.W_K, a.W_V)):
                ent.append((w, (H, D, dh, 1), (D * dh, dh, 1, 0), 0, j * hd * D, 0))
                ent.append((w, (1, D, H, dh), (0, dh, D * dh, 1), 1, j * hd, 3 * hd))
            specs = tuple(((H, D, dh), (dh * D, 1, D), j * hd * D) for j in range(3))
            sites.append(dict(key=("qkv", i), n_in=D, n_out=3 * hd, n_out_pad=3 * hd, params=(a.W_Q, a.W_K, a.W_V), entries=ent,
                              specs=specs))
            # [in][out] row-major matrices: filter view [K=out, C=in] with strides (1, out)
            for key, w, n_in, n_out, size, stride in (
                    (("o", i), a.W_O, hd, D, (H, dh, D), (dh, 1, hd)),
                    (("in", i), m.W_in, D, F_, (D, F_), (1, D)),
                    (("out", i), m.W_out, F_, D, (F_, D), (1, F_))):
                ent = [(w, (n_out, n_in, 1, 1), (1, n_out, 0, 0), 0, 0, 0), (w, (n_out, n_in, 1, 1), (1, n_out, 0, 0), 1, 0, 0)]
                sites.append(dict(key=key, n_in=n_in, n_out=n_out, n_out_pad=n_out, params=(w,), entries=ent,
                                  specs=((size, stride, 0),)))
        Vp = V + (-V) % 64
        w = self.unembed.W_U
        ent = [(w, (V, D, 1, 1), (1, V, 0, 0), 0, 0, 0), (w, (V, D, 1, 1), (1, V, 0, 0), 1, 0, Vp)]
        sites.append(dict(key=("unembed",), n_in=D, n_out=V, n_out_pad=Vp, params=(w,), entries=ent, specs=(((D, V), (1, D), 0),)))
        return sites

    def _shadow_usable(self):
        c = self.cfg
        return (self.act_dtype == torch.bfloat16 and ops.USE_TENSOR_CORES and c.d_model % 64 == 0 and c.d_mlp % 64 == 0
                and (c.n_heads * c.d_head) % 64 == 0 and self.unembed.W_U.is_cuda)

    def _build_filter_shadow(self):
        from .._lib import IIT_BF16, Relayout
        sites = self._linear_sites()
        dev = self.unembed.W_U.device
        n_ent = sum(len(st["entries"]) for st in sites)
        table = (Relayout * n_ent)()
        bufs, k, biggest = {}, 0, 1
        for st in sites:
            oi = torch.zeros(st["n_out_pad"] * st["n_in"], dtype=torch.bfloat16, device=dev)
            io = torch.zeros(st["n_in"] * st["n_out_pad"], dtype=torch.bfloat16, device=dev)
            bufs[st["key"]] = (oi, io, st)
            for w, dims, strides, to_crsk, off, pitch in st["entries"]:
                e = table[k]; k += 1
                e.w, e.out = w.data_ptr(), (io if to_crsk else oi).data_ptr() + 2 * off
                for d in range(4):
                    e.stride[d] = strides[d]
                e.k, e.c, e.r, e.s = dims
                e.to_crsk, e.dtype, e.out_pitch = to_crsk, IIT_BF16, pitch
                biggest = max(biggest, dims[0] * dims[1] * dims[2] * dims[3])
        raw = torch.frombuffer(bytearray(bytes(table)), dtype=torch.uint8).clone()
        params = [p for st in sites for p in st["params"]]
        self._shadow = {"bufs": bufs, "table": raw.to(dev), "entries": n_ent, "max": biggest, "params": params,
                        "ptrs": [p.data_ptr() for p in params], "versions": None}

    def refresh_filter_shadow(self):
        if not self._shadow_usable():
            return
        sh = getattr(self, "_shadow", None)
        if sh is None or sh["ptrs"] != [p.data_ptr() for p in sh["params"]]:
            if ops.capturing():
                return                       # never (re)allocate inside a capture: forward falls back to per-GEMM copies
            self._build_filter_shadow()
            sh = self._shadow
        ops.filter_relayout_batched(sh["table"], sh["entries"], sh["max"])
        sh["versions"] = [p._version for p in sh["params"]]
        # the unembedding bias padded to the GEMM's width (fp32: the epilogue adds it to the accumulator)
        V = self.cfg.d_vocab_out
        bu = sh.get("b_U_pad")
        if bu is None:
            bu = sh["b_U_pad"] = torch.zeros(V + (-V) % 64, dtype=torch.float32, device=self.unembed.b_U.device)
        with torch.no_grad():
            bu[:V].copy_(self.unembed.b_U)
        sh["b_U_version"] = self.unembed.b_U._version

    def _filter_shadow_stale(self):
        if not self._shadow_usable():
            return False
        sh = getattr(self, "_shadow", None)
        return (sh is None or sh["versions"] != [p._version for p in sh["params"]]
                or sh["ptrs"] != [p.data_ptr() for p in sh["params"]] or sh.get("b_U_version") != self.unembed.b_U._version)

    # ---- fused epilogues (bf16 prepared path): bias / residual / gelu_new in the GEMM, GELU backward in the output map's dgrad
    def _qkv_bias(self, a):
        """b_Q | b_K | b_V as one fp32 [3*H*dh] vector: a view when the three parameters are adjacent in memory (they are
        under FlatAdam, whose flat buffer follows parameter order), a concatenation otherwise."""
        n = a.b_Q.numel()
        if (a.b_Q.is_contiguous() and a.b_K.data_ptr() == a.b_Q.data_ptr() + 4 * n and a.b_V.data_ptr() == a.b_Q.data_ptr() + 8 * n
                and a.b_Q.untyped_storage().data_ptr() == a.b_V.untyped_storage().data_ptr()):
            return torch.as_strided(a.b_Q.detach(), (3 * n,), (1,), a.b_Q.storage_offset())
        return torch.cat([b.detach().reshape(-1) for b in (a.b_Q, a.b_K, a.b_V)])

    @staticmethod
    def _direct_bias_grad(params):
        """The optimizer's fp32 gradient buffer behind one bias parameter -- or behind several that lie back to back in it
        (b_Q | b_K | b_V under FlatAdam) -- when backward kernels may add into it (ops.direct_grad); None otherwise."""
        if not torch.is_grad_enabled():
            return None
        gs = [ops.direct_grad(p) for p in params]
        if any(g is None or not g.is_contiguous() for g in gs):
            return None
        n = gs[0].numel()
        for j, g in enumerate(gs):
            if g.numel() != n or g.data_ptr() != gs[0].data_ptr() + 4 * n * j:
                return None
        return gs[0].reshape(-1) if len(gs) == 1 else torch.as_strided(gs[0], (n * len(gs),), (1,), gs[0].storage_offset())

    def _epi(self, key, x, bias, bias_specs, bias_params, resid=None, act=0, link_out=None, link_in=None, direct_bias=True,
             res_slot=None):
        if direct_bias and act == 0:
            buf = self._direct_bias_grad(bias_params)
            if buf is not None:
                bias_specs = (*bias_specs, ("direct", buf))
        oi, io, st = self._shadow["bufs"][key]
        specs = (ops.wgrad_direct_maps(st["specs"], st["params"], st) if DIRECT_WEIGHT_GRADS else None) or st["specs"]
        return ops.linear_epi(x, oi, io, st["n_in"], st["n_out_pad"], specs, st["params"], bias=bias, bias_specs=bias_specs,
                              bias_params=bias_params, resid=resid, act=act, link_out=link_out, link_in=link_in, res_slot=res_slot)

    def _mlp_link(self, li):
        sites = self.__dict__.setdefault("_mlp_sites", {})
        ent = sites.get(li)
        dev = self.unembed.W_U.device
        if ent is None or ent[0].device != dev:
            ent = sites[li] = (torch.zeros((2, self.cfg.d_mlp), dtype=torch.float64, device=dev), {"dirty": False})
        return ops.GeluLink(ent[0], ent[1])

    def _linear(self, key, x):
        """x @ W of site `key` from the prepared operands, or None when they are not in use."""
        sh = getattr(self, "_shadow", None)
        if sh is None or sh["versions"] is None:
            return None
        oi, io, st = sh["bufs"][key]
        specs = (ops.wgrad_direct_maps(st["specs"], st["params"], st) if DIRECT_WEIGHT_GRADS else None) or st["specs"]
        return ops.linear_prepared(x, oi, io, st["n_in"], st["n_out_pad"], specs, st["params"])

    # ------------------------------------------------------------------------------------------
    def _check_hooks(self):
        for n in self._unsupported:
            if self.hook_dict[n].has_hooks():
                raise NotImplementedError(f"hook point {n} lives inside a fused kernel; hook hook_z / hook_post / "
                                          "the residual-stream hook points instead")

    def forward(self, tokens, last_pos_only=False, return_type="logits", return_padded=False):
        """`return_padded` (with last_pos_only): the last-position logits as the unembedding GEMM wrote them, [B, V rounded up
        to 64 columns] -- the caller scores columns < d_vocab_out only (ops.cross_entropy(..., classes=V)) and no slice sits
        in the autograd graph (its backward would rebuild the padded gradient with a fill and a copy)."""
        self._check_hooks()
        if self._shadow_usable():
            if self._filter_shadow_stale() and not ops.capturing():
                self.refresh_filter_shadow()
            use_shadow = not self._filter_shadow_stale()
        else:
            use_shadow = False
        c = self.cfg
        B, S = tokens.shape
        D, H, dh = c.d_model, c.n_heads, c.d_head
        T = B * S
        resid = ops.embed(tokens, self.embed.W_E, self.pos_embed.W_pos, self.act_dtype).view(T, D)
        blocks_seen = []
        for blk in self.blocks:
            a, m = blk.attn, blk.mlp
            resid = blk.hook_resid_pre(resid.view(B, S, D)).reshape(T, D)
            fused = use_shadow and not (blk.hook_attn_out.has_hooks() or blk.hook_mlp_out.has_hooks() or m.hook_pre.has_hooks())
            # the residual stream feeds a LayerNorm and a skip connection: their two gradients meet in a GradSlot
            slot_a = ops.GradSlot() if (fused and torch.is_grad_enabled() and resid.requires_grad) else None
            x = ops.lnpre(resid, c.eps, slot_a)
            if fused:
                # every element-wise neighbour of the four linear maps runs in their GEMM epilogues
                li = len(blocks_seen); blocks_seen.append(blk)
                hd = H * dh
                qkv = self._epi(("qkv", li), x, self._qkv_bias(a), tuple(((H, dh), (dh, 1), j * hd) for j in range(3)),
                                (a.b_Q, a.b_K, a.b_V)).view(B, S, 3, H, dh)
                z = a.hook_z(ops.attention(qkv))
                resid = self._epi(("o", li), z.reshape(T, hd), a.b_O.detach(), (((D,), (1,), 0),), (a.b_O,), resid=resid, res_slot=slot_a)
                resid = blk.hook_resid_mid(resid.view(B, S, D)).reshape(T, D)
                slot_m = ops.GradSlot() if (torch.is_grad_enabled() and resid.requires_grad) else None
                x = ops.lnpre(resid, c.eps, slot_m)
                if FUSE_GELU_IN_GEMM:      # measured slower: the tiled kernel's eight epilogue warps are ALU-bound on the GELU
                    link = self._mlp_link(li) if (torch.is_grad_enabled() and m.hook_post.only_patch_hooks()) else None
                    post = self._epi(("in", li), x, m.b_in.detach(), (((c.d_mlp,), (1,), 0),), (m.b_in,), act=1, link_out=link)
                    post = post.view(B, S, c.d_mlp)
                else:
                    # the MLP's input map stays a plain GEMM; bias + gelu_new read its result once and write `post` only
                    # (the backward recomputes h + bias and takes the bias gradient in the same pass)
                    link = None
                    post = ops.bias_gelu_lean(self._linear(("in", li), x), m.b_in, (B, S, c.d_mlp), self._direct_bias_grad((m.b_in,)))
                post = m.hook_post(post).reshape(T, -1)
                resid = self._epi(("out", li), post, m.b_out.detach(), (((D,), (1,), 0),), (m.b_out,), resid=resid, link_in=link,
                                  res_slot=slot_m)
                resid = blk.hook_resid_post(resid.view(B, S, D)).reshape(T, D)
                continue
            # fused QKV projection: [D, (3,H,dh)] assembled from the per-head parameters (tiny; autograd routes grads)
            b_qkv = torch.cat([b.reshape(-1) for b in (a.b_Q, a.b_K, a.b_V)])
            if use_shadow:
                li = len(blocks_seen); blocks_seen.append(blk)
                qkv_lin = self._linear(("qkv", li), x)
            else:
                w_qkv = torch.cat([w.permute(1, 0, 2).reshape(D, H * dh) for w in (a.W_Q, a.W_K, a.W_V)], dim=1)
                qkv_lin = ops.linear_tokens(x, w_qkv)
            qkv = ops.bias_add(qkv_lin, b_qkv).view(B, S, 3, H, dh)
            z = a.hook_z(ops.attention(qkv))
            zr = z.reshape(T, H * dh)
            proj = self._linear(("o", li), zr) if use_shadow else ops.linear_tokens(zr, a.W_O.reshape(H * dh, D))
            if blk.hook_attn_out.has_hooks():
                attn_out = blk.hook_attn_out(ops.bias_add(proj, a.b_O).view(B, S, D)).reshape(T, D)
                resid = ops.bias_add(attn_out, None, resid)
            else:
                resid = ops.bias_add(proj, a.b_O, resid)
            resid = blk.hook_resid_mid(resid.view(B, S, D)).reshape(T, D)
            x = ops.lnpre(resid, c.eps)
            h = self._linear(("in", li), x) if use_shadow else ops.linear_tokens(x, m.W_in)
            if m.hook_pre.has_hooks():  # observers only: the activation below is computed from the unhooked value
                pre = ops.bias_add(h, m.b_in).view(B, S, -1)
                if m.hook_pre(pre) is not pre:
                    raise NotImplementedError("mlp.hook_pre is read-only here; hook mlp.hook_post to intervene")
            post = m.hook_post(ops.bias_gelu(h, m.b_in, (B, S, c.d_mlp))).reshape(T, -1)
            out = self._linear(("out", li), post) if use_shadow else ops.linear_tokens(post, m.W_out)
            if blk.hook_mlp_out.has_hooks():
                mlp_out = blk.hook_mlp_out(ops.bias_add(out, m.b_out).view(B, S, D)).reshape(T, D)
                resid = ops.bias_add(mlp_out, None, resid)
            else:
                resid = ops.bias_add(out, m.b_out, resid)
            resid = blk.hook_resid_post(resid.view(B, S, D)).reshape(T, D)
        W_U, b_U, V = self.unembed.W_U, self.unembed.b_U, c.d_vocab_out
        pad = (-V) % 64
        if use_shadow:
            xin = resid.view(B, S, D)[:, -1, :].contiguous() if last_pos_only else resid
            # (the padded width is wider than b_U's gradient: its column sums go through autograd)
            logits = self._epi(("unembed",), ops.lnpre(xin, c.eps), self._shadow["b_U_pad"], (((V,), (1,), 0),), (b_U,), direct_bias=False)
            if last_pos_only and return_padded:
                return logits
            logits = logits[:, :V]
            return logits if last_pos_only else logits.reshape(B, S, -1)
        if pad and self.act_dtype == torch.bfloat16 and ops.USE_TENSOR_CORES:
            # the tensor-core GEMMs want widths that are multiples of 64: unembed into V + pad columns (zero weights and
            # bias) and slice them off again; autograd routes the gradient (50257 -> 50304 for GPT-2's vocabulary)
            W_U = torch.nn.functional.pad(W_U, (0, pad))
            b_U = torch.nn.functional.pad(b_U, (0, pad))
        if last_pos_only:
            last = resid.view(B, S, D)[:, -1, :].contiguous()
            return ops.bias_add(ops.linear_tokens(ops.lnpre(last, c.eps), W_U), b_U)[:, :V]
        logits = ops.bias_add(ops.linear_tokens(ops.lnpre(resid, c.eps), W_U), b_U)[:, :V]
        return logits.reshape(B, S, -1)
