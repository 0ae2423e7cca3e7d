"""TEST INFRASTRUCTURE (oracle) -- restatement of transformer_lens==1.12.1 `HookedTransformer`.

[TL-knowledge] Minimal GPT-2-style decoder exactly as the reference configures it
(/root/reference/train_ioi.py:30-34 with iit/tasks/ioi/__init__.py:8-17): standard learned
positional embedding, pre-norm blocks, LN or LNPre (folded layernorm), causal attention with
per-head W_Q/W_K/W_V [H, d_model, d_head] and W_O [H, d_head, d_model], MLP with
gelu_new / gelu / relu, un-tied unembed. No tokenizer, no from_pretrained (no network).
Never imported by the product path.
"""
from __future__ import annotations

import math
from types import SimpleNamespace

import numpy as np
import torch
import torch.nn as nn
import torch.nn.functional as F

from .hook_points import HookedRootModule, HookPoint  # noqa: F401  (HookPoint re-exported: eval_ablations.py:9)


class HookedTransformerConfig(SimpleNamespace):
    DEFAULTS = dict(
        n_layers=None, d_model=None, n_ctx=None, d_head=None, model_name="custom", n_heads=-1,
        d_mlp=None, act_fn=None, d_vocab=-1, eps=1e-5, use_attn_result=False, use_attn_scale=True,
        use_split_qkv_input=False, use_hook_mlp_in=False, use_local_attn=False,
        original_architecture=None, from_checkpoint=False, tokenizer_name=None,
        normalization_type="LN", device=None, n_devices=1, attention_dir="causal", attn_only=False,
        seed=None, initializer_range=-1.0, init_weights=True, positional_embedding_type="standard",
        d_vocab_out=-1, default_prepend_bos=True, dtype=torch.float32,
    )

    def __init__(self, **kw):
        d = dict(self.DEFAULTS)
        for k, v in kw.items():
            d[k] = v  # unknown keys (e.g. from a gpt2 cfg dict) are carried along
        super().__init__(**d)
        if self.n_heads == -1:
            self.n_heads = self.d_model // self.d_head
        if self.d_vocab_out == -1:
            self.d_vocab_out = self.d_vocab
        if not self.attn_only and self.d_mlp is None:
            self.d_mlp = self.d_model * 4
        if self.initializer_range < 0:
            self.initializer_range = 0.8 / math.sqrt(self.d_model)
        if self.device is None:
            self.device = "cuda" if torch.cuda.is_available() else "cpu"
        self.attn_scale = math.sqrt(self.d_head) if self.use_attn_scale else 1.0

    def to_dict(self):
        return dict(self.__dict__)


def gelu_new(x):
    return 0.5 * x * (1.0 + torch.tanh(math.sqrt(2.0 / math.pi) * (x + 0.044715 * torch.pow(x, 3.0))))


class LayerNormPre(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.eps = cfg.eps
        self.hook_scale = HookPoint()
        self.hook_normalized = HookPoint()

    def forward(self, x):
        x = x - x.mean(-1, keepdim=True)
        scale = self.hook_scale((x.pow(2).mean(-1, keepdim=True) + self.eps).sqrt())
        return self.hook_normalized(x / scale)


class LayerNorm(nn.Module):
    def __init__(self, cfg, length=None):
        super().__init__()
        self.eps = cfg.eps
        length = cfg.d_model if length is None else length
        self.w = nn.Parameter(torch.ones(length))
        self.b = nn.Parameter(torch.zeros(length))
        self.hook_scale = HookPoint()
        self.hook_normalized = HookPoint()

    def forward(self, x):
        x = x - x.mean(-1, keepdim=True)
        scale = self.hook_scale((x.pow(2).mean(-1, keepdim=True) + self.eps).sqrt())
        x = x / scale
        return self.hook_normalized(x * self.w + self.b)


class Identity(nn.Module):
    def forward(self, x):
        return x


def make_norm(cfg):
    if cfg.normalization_type == "LN":
        return LayerNorm(cfg)
    if cfg.normalization_type == "LNPre":
        return LayerNormPre(cfg)
    if cfg.normalization_type is None:
        return Identity()
    raise ValueError(cfg.normalization_type)


class Embed(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.W_E = nn.Parameter(torch.empty(cfg.d_vocab, cfg.d_model))

    def forward(self, tokens):
        return self.W_E[tokens, :]


class PosEmbed(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.W_pos = nn.Parameter(torch.empty(cfg.n_ctx, cfg.d_model))

    def forward(self, tokens):
        b, s = tokens.shape
        return self.W_pos[:s, :].unsqueeze(0).expand(b, s, -1)


class Unembed(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.W_U = nn.Parameter(torch.empty(cfg.d_model, cfg.d_vocab_out))
        self.b_U = nn.Parameter(torch.zeros(cfg.d_vocab_out))

    def forward(self, resid):
        return resid @ self.W_U + self.b_U


class Attention(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.cfg = cfg
        H, D, dh = cfg.n_heads, cfg.d_model, cfg.d_head
        self.W_Q = nn.Parameter(torch.empty(H, D, dh))
        self.W_K = nn.Parameter(torch.empty(H, D, dh))
        self.W_V = nn.Parameter(torch.empty(H, D, dh))
        self.W_O = nn.Parameter(torch.empty(H, dh, D))
        self.b_Q = nn.Parameter(torch.zeros(H, dh))
        self.b_K = nn.Parameter(torch.zeros(H, dh))
        self.b_V = nn.Parameter(torch.zeros(H, dh))
        self.b_O = nn.Parameter(torch.zeros(D))
        self.register_buffer("mask", torch.tril(torch.ones(cfg.n_ctx, cfg.n_ctx).bool()), persistent=True)
        self.register_buffer("IGNORE", torch.tensor(-torch.inf), persistent=True)
        self.hook_k = HookPoint()
        self.hook_q = HookPoint()
        self.hook_v = HookPoint()
        self.hook_z = HookPoint()
        self.hook_attn_scores = HookPoint()
        self.hook_pattern = HookPoint()
        self.hook_result = HookPoint()

    def forward(self, x):
        q = self.hook_q(torch.einsum("bpm,hmd->bphd", x, self.W_Q) + self.b_Q)
        k = self.hook_k(torch.einsum("bpm,hmd->bphd", x, self.W_K) + self.b_K)
        v = self.hook_v(torch.einsum("bpm,hmd->bphd", x, self.W_V) + self.b_V)
        scores = torch.einsum("bqhd,bkhd->bhqk", q, k) / self.cfg.attn_scale
        if self.cfg.attention_dir == "causal":
            s = scores.shape[-1]
            scores = torch.where(self.mask[:s, :s], scores, self.IGNORE)
        scores = self.hook_attn_scores(scores)
        pattern = self.hook_pattern(F.softmax(scores, dim=-1))
        pattern = torch.where(torch.isnan(pattern), torch.zeros_like(pattern), pattern)
        z = self.hook_z(torch.einsum("bkhd,bhqk->bqhd", v, pattern))
        if not self.cfg.use_attn_result:
            return torch.einsum("bqhd,hdm->bqm", z, self.W_O) + self.b_O
        result = self.hook_result(torch.einsum("bqhd,hdm->bqhm", z, self.W_O))
        return result.sum(dim=2) + self.b_O


class MLP(nn.Module):
    def __init__(self, cfg):
        super().__init__()
        self.cfg = cfg
        self.W_in = nn.Parameter(torch.empty(cfg.d_model, cfg.d_mlp))
        self.b_in = nn.Parameter(torch.zeros(cfg.d_mlp))
        self.W_out = nn.Parameter(torch.empty(cfg.d_mlp, cfg.d_model))
        self.b_out = nn.Parameter(torch.zeros(cfg.d_model))
        self.hook_pre = HookPoint()
        self.hook_post = HookPoint()
        self.act_fn = {"relu": F.relu, "gelu": F.gelu, "gelu_new": gelu_new, "silu": F.silu}[cfg.act_fn]

    def forward(self, x):
        pre = self.hook_pre(x @ self.W_in + self.b_in)
        post = self.hook_post(self.act_fn(pre))
        return post @ self.W_out + self.b_out


class TransformerBlock(nn.Module):
    def __init__(self, cfg, block_index):
        super().__init__()
        self.cfg = cfg
        self.ln1 = make_norm(cfg)
        if not cfg.attn_only:
            self.ln2 = make_norm(cfg)
        self.attn = Attention(cfg)
        if not cfg.attn_only:
            self.mlp = MLP(cfg)
        self.hook_attn_in = HookPoint()
        self.hook_q_input = HookPoint()
        self.hook_k_input = HookPoint()
        self.hook_v_input = HookPoint()
        self.hook_mlp_in = HookPoint()
        self.hook_attn_out = HookPoint()
        self.hook_mlp_out = HookPoint()
        self.hook_resid_pre = HookPoint()
        if not cfg.attn_only:
            self.hook_resid_mid = HookPoint()
        self.hook_resid_post = HookPoint()

    def forward(self, resid_pre):
        resid_pre = self.hook_resid_pre(resid_pre)
        attn_out = self.hook_attn_out(self.attn(self.ln1(resid_pre)))
        if self.cfg.attn_only:
            return self.hook_resid_post(resid_pre + attn_out)
        resid_mid = self.hook_resid_mid(resid_pre + attn_out)
        mlp_out = self.hook_mlp_out(self.mlp(self.ln2(resid_mid)))
        return self.hook_resid_post(resid_mid + mlp_out)


class HookedTransformer(HookedRootModule):
    def __init__(self, cfg, tokenizer=None, move_to_device=True, default_padding_side="right"):
        super().__init__()
        if isinstance(cfg, dict):
            cfg = HookedTransformerConfig(**cfg)
        self.cfg = cfg
        self.tokenizer = tokenizer
        self.embed = Embed(cfg)
        self.hook_embed = HookPoint()
        self.pos_embed = PosEmbed(cfg)
        self.hook_pos_embed = HookPoint()
        self.blocks = nn.ModuleList([TransformerBlock(cfg, i) for i in range(cfg.n_layers)])
        self.ln_final = make_norm(cfg)
        self.unembed = Unembed(cfg)
        if cfg.init_weights:
            self.init_weights()
        self.setup()

    def init_weights(self):
        if self.cfg.seed is not None:
            torch.manual_seed(self.cfg.seed)
        for name, param in self.named_parameters():
            if "W_" in name:
                nn.init.normal_(param, std=self.cfg.initializer_range)

    def forward(self, tokens, return_type="logits"):
        embed = self.hook_embed(self.embed(tokens))
        pos_embed = self.hook_pos_embed(self.pos_embed(tokens))
        residual = embed + pos_embed
        for block in self.blocks:
            residual = block(residual)
        residual = self.ln_final(residual)
        return self.unembed(residual)
