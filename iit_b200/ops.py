"""Torch-facing operators over the C-ABI (iit_b200/_lib.py -> libiit_b200.so).

PyTorch is plumbing here: it owns device memory, streams and the autograd tape; every operator's
arithmetic runs in the hand-written sm_100a kernels. There is no CPU / eager fallback: tensors that
are not on a CUDA device raise.

Layout convention for image activations: logical shape [N, C, H, W] with channels-last strides
(physical NHWC), so the reference-facing view is NCHW while C is the contiguous dim in memory.
"""
from __future__ import annotations

import ctypes as C
import weakref
from typing import Optional

import torch

from . import _lib
from ._lib import Box, Conv, call

_DT = {torch.float32: _lib.IIT_F32, torch.bfloat16: _lib.IIT_BF16}
# bf16 convolutions that qualify run on the tcgen05 tensor-core kernels; the CUDA-core kernels remain the
# exact-fp32 path and the catch-all. (A switch for A/B measurements, not a fallback: both are this library.)
USE_TENSOR_CORES = True


def capturing() -> bool:
    """True while the current CUDA stream is being captured into a graph."""
    return torch.cuda.is_available() and torch.cuda.is_current_stream_capturing()


def _require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("iit_b200 kernels need CUDA tensors (no CPU fallback exists); got a tensor on "
                               f"{t.device}")


def _p(t: Optional[torch.Tensor]):
    return None if t is None else C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _dt(t: torch.Tensor) -> int:
    try:
        return _DT[t.dtype]
    except KeyError:
        raise TypeError(f"unsupported dtype {t.dtype}; expected float32 or bfloat16") from None


def _is_cl(t: torch.Tensor) -> bool:
    return t.dim() == 4 and t.is_contiguous(memory_format=torch.channels_last)


def _cl(t: torch.Tensor) -> torch.Tensor:
    """channels-last contiguous (no copy when already so)."""
    return t if _is_cl(t) else t.contiguous(memory_format=torch.channels_last)


def empty_cl(n, c, h, w, dtype, device):
    return torch.empty((n, c, h, w), dtype=dtype, device=device, memory_format=torch.channels_last)


# =================================================================================================
# (a) patch
# =================================================================================================
def make_box(index, dst: torch.Tensor, src: Optional[torch.Tensor]) -> Optional[Box]:
    """Box descriptor for `dst[index] = src[index]` from a TorchIndex and the tensors' strides."""
    spans = index.box(tuple(dst.shape))
    if spans is None:
        return None
    if dst.dim() > _lib.IIT_MAX_RANK:
        raise ValueError(f"rank {dst.dim()} exceeds {_lib.IIT_MAX_RANK}")
    b = Box()
    b.rank = dst.dim()
    b.elem_bytes = dst.element_size()
    for d, (start, extent) in enumerate(spans):
        b.start[d], b.extent[d] = start, extent
        b.dst_stride[d] = dst.stride(d)
        b.src_stride[d] = src.stride(d) if src is not None else dst.stride(d)
    return b


def box_bytes(index, t: torch.Tensor) -> int:
    spans = index.box(tuple(t.shape))
    n = 1
    for _, e in spans:
        n *= e
    return n * t.element_size()


def patch_(dst: torch.Tensor, src: torch.Tensor, index) -> torch.Tensor:
    """In place, no autograd: dst[index] = src[index]."""
    _require_cuda(dst, src)
    if dst.shape != src.shape or dst.dtype != src.dtype:
        raise ValueError(f"patch: shape/dtype mismatch {tuple(dst.shape)} {dst.dtype} vs {tuple(src.shape)} {src.dtype}")
    box = make_box(index, dst, src)
    if box is None:  # list-of-ints index: no box form -> advanced-indexing gather/scatter by torch
        dst[index.as_index] = src[index.as_index]
        return dst
    call("iit_patch_fwd", _p(dst), _p(src), C.byref(box), _stream())
    return dst


def zero_box_(grad: torch.Tensor, index) -> torch.Tensor:
    _require_cuda(grad)
    box = make_box(index, grad, None)
    if box is None:
        grad[index.as_index] = 0
        return grad
    call("iit_patch_bwd", _p(grad), C.byref(box), _stream())
    return grad


def capture_box(src: torch.Tensor, index) -> torch.Tensor:
    """Compact copy of src[index] (box order), for caches that must outlive the source tensor."""
    _require_cuda(src)
    spans = index.box(tuple(src.shape))
    if spans is None:
        return src[index.as_index].contiguous()
    out = torch.empty([e for _, e in spans], dtype=src.dtype, device=src.device)
    box = make_box(index, src, src)
    call("iit_capture_box", _p(out), _p(src), C.byref(box), _stream())
    return out


class _PatchFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, act, cached, index, inplace):
        ctx.index = index
        ctx.whole = index.is_whole()
        if inplace and not (act.is_leaf and act.requires_grad):
            out = act
            ctx.mark_dirty(act)
        else:
            out = act.clone(memory_format=torch.preserve_format)
        patch_(out, cached, index)
        return out

    @staticmethod
    def backward(ctx, g):
        if ctx.whole:
            return None, None, None, None  # every element was overwritten by a constant
        g = g.clone(memory_format=torch.preserve_format)
        zero_box_(g, ctx.index)
        return g, None, None, None


def patch(act: torch.Tensor, cached: torch.Tensor, index, inplace: bool = True) -> torch.Tensor:
    """`out = act.clone(); out[index] = cached[index]` with the CopySlices gradient (zero inside the
    box, identity outside). With inplace=True `act` itself is overwritten (legal for a hook-point
    value that no earlier op saved for backward)."""
    return _PatchFn.apply(act, cached.detach(), index, inplace)


# =================================================================================================
# (c) high-level models + losses
# =================================================================================================
def hl_pvr(base_vars: torch.Tensor, src_vars: Optional[torch.Tensor], var: int) -> torch.Tensor:
    _require_cuda(base_vars, src_vars)
    bv = base_vars.contiguous()
    sv = src_vars.contiguous() if src_vars is not None else None
    if bv.dtype != torch.int64 or bv.dim() != 2 or bv.shape[1] != 4:
        raise ValueError("hl_pvr: intermediate vars must be int64 [B,4]")
    out = torch.empty(bv.shape[0], dtype=torch.int64, device=bv.device)
    # a top-left variable outside [-10, 9] (bad data, a user hook on hook_tl) sets the device error flag and yields -1;
    # raise_on_index_error() turns it into the IndexError the reference's indexing raises
    call("iit_hl_pvr", _p(bv), _p(sv), int(var), bv.shape[0], _p(out), _p(_err_flag(bv.device)), _stream())
    return out


def hl_ioi_last_argmax(base_tok, src_tok, node: Optional[str], names, d_vocab: int) -> torch.Tensor:
    _require_cuda(base_tok, src_tok, names)
    bt = base_tok.contiguous()
    st = src_tok.contiguous() if src_tok is not None else None
    out = torch.empty(bt.shape[0], dtype=torch.int64, device=bt.device)
    call("iit_hl_ioi_last_argmax", _p(bt), _p(st), _lib.IOI_NODE_IDS[node], _p(names), int(names.numel()),
         bt.shape[0], bt.shape[1], int(d_vocab), _p(out), _stream())
    return out


def hl_ioi_dense(base_tok, src_tok, node: Optional[str], names, d_vocab: int) -> torch.Tensor:
    _require_cuda(base_tok, src_tok, names)
    bt = base_tok.contiguous()
    st = src_tok.contiguous() if src_tok is not None else None
    out = torch.empty((bt.shape[0], bt.shape[1], d_vocab), dtype=torch.float32, device=bt.device)
    call("iit_hl_ioi_dense", _p(bt), _p(st), _lib.IOI_NODE_IDS[node], _p(names), int(names.numel()),
         bt.shape[0], bt.shape[1], int(d_vocab), _p(out), _stream())
    return out


def _rows_view(logits: torch.Tensor):
    """[rows, classes] with unit class stride and a single row stride, without copying if possible."""
    if logits.dim() != 2:
        raise ValueError("expected 2-D logits [rows, classes]")
    if logits.stride(1) != 1:
        logits = logits.contiguous()
    return logits, logits.stride(0)


_tickets: dict = {}


def _ticket(device) -> torch.Tensor:
    """One zeroed uint32 per device for iit_hl_ce_fwd's last-block reduction (the kernel leaves it zero)."""
    dev = torch.device(device)
    if dev.type == "cuda" and dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    t = _tickets.get(dev)
    if t is None:
        t = _tickets[dev] = torch.zeros(1, dtype=torch.int32, device=dev)
    return t


class _HLCEFn(torch.autograd.Function):
    """Mean cross-entropy of `logits[:, :classes]` against a class index per row that comes from a high-level model, all in
    one launch (iit_hl_ce_fwd): `hl` = None (targets given), ("pvr", base_vars, src_vars, var) or
    ("ioi", base_tok, src_tok, node, names, d_vocab). `logits` may be wider than `classes` (padded columns): the gradient is
    then as wide, zero in the padding. Returns (loss, target)."""

    @staticmethod
    def forward(ctx, logits, target, classes, hl):
        _require_cuda(logits, target)
        lg, rs = _rows_view(logits)
        rows = lg.shape[0]
        classes = lg.shape[1] if classes is None else int(classes)
        spec = _lib.HlSpec()
        keep = []
        if hl is None:
            spec.mode = 0
            tgt = target.contiguous()
        else:
            tgt = torch.empty(rows, dtype=torch.int64, device=lg.device)
            if hl[0] == "pvr":
                _, bv, sv, var = hl
                bv = bv.contiguous()
                sv = sv.contiguous() if sv is not None else None
                if bv.dtype != torch.int64 or bv.dim() != 2 or bv.shape[1] != 4 or bv.shape[0] != rows:
                    raise ValueError("PVR HL: intermediate vars must be int64 [rows,4]")
                spec.mode, spec.node, spec.base, spec.src = 1, int(var), bv.data_ptr(), (sv.data_ptr() if sv is not None else None)
                keep += [bv, sv]
            else:
                _, bt, st_, node, names, d_vocab = hl
                bt = bt.contiguous()
                st_ = st_.contiguous() if st_ is not None else None
                if bt.shape[0] != rows:
                    raise ValueError("IOI HL: one prompt per row of logits expected")
                spec.mode, spec.node, spec.base = 2, _lib.IOI_NODE_IDS[node], bt.data_ptr()
                spec.src = st_.data_ptr() if st_ is not None else None
                spec.names, spec.n_names, spec.seq, spec.d_vocab = names.data_ptr(), int(names.numel()), int(bt.shape[1]), int(d_vocab)
                keep += [bt, st_, names]
        lse = torch.empty(rows, dtype=torch.float32, device=lg.device)
        nll = torch.empty(rows, dtype=torch.float32, device=lg.device)
        loss = torch.empty((), dtype=torch.float32, device=lg.device)
        call("iit_hl_ce_fwd", _p(lg), _dt(lg), rows, classes, rs, C.byref(spec), _p(tgt), _p(lse), _p(nll), _p(loss),
             _p(_ticket(lg.device)), _p(_err_flag(lg.device)), _stream())
        ctx.save_for_backward(lg, tgt, lse)
        ctx.classes = classes
        ctx.mark_non_differentiable(tgt)
        return loss, tgt

    @staticmethod
    def backward(ctx, g, _g_tgt):
        lg, tgt, lse = ctx.saved_tensors
        rows, width = lg.shape
        d = torch.empty((rows, width), dtype=lg.dtype, device=lg.device)
        if width > ctx.classes and not (lg.dtype == torch.bfloat16 and width % 8 == 0):
            d[:, ctx.classes:] = 0          # the vector path writes the padding columns itself
        g = g.to(torch.float32).contiguous()
        call("iit_ce_index_bwd", _p(lg), _dt(lg), rows, ctx.classes, lg.stride(0), _p(tgt), _p(lse), _p(g), _p(d), width,
             _stream())
        return d, None, None, None


class _CEIndexFn:
    @staticmethod
    def apply(logits, target):
        return _HLCEFn.apply(logits, target, None, None)[0]


def hl_cross_entropy(logits, hl, classes=None):
    """(loss, hl_target): see _HLCEFn."""
    return _HLCEFn.apply(logits, None, classes, hl)


class _CEProbFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, logits, prob):
        _require_cuda(logits, prob)
        lg, rs = _rows_view(logits)
        pr = prob if (prob.dtype == torch.float32 and prob.stride(1) == 1) else prob.float().contiguous()
        rows, classes = lg.shape
        lse = torch.empty(rows, dtype=torch.float32, device=lg.device)
        psum = torch.empty(2 * rows, dtype=torch.float32, device=lg.device)  # [psum | per-row nll]
        loss = torch.zeros((), dtype=torch.float32, device=lg.device)
        call("iit_ce_prob_fwd", _p(lg), _dt(lg), rows, classes, rs, _p(pr), pr.stride(0), _p(lse), _p(psum), _p(loss),
             _stream())
        ctx.save_for_backward(lg, pr, lse, psum)
        return loss

    @staticmethod
    def backward(ctx, g):
        lg, pr, lse, psum = ctx.saved_tensors
        rows, classes = lg.shape
        d = torch.empty((rows, classes), dtype=lg.dtype, device=lg.device)
        g = g.to(torch.float32).contiguous()
        call("iit_ce_prob_bwd", _p(lg), _dt(lg), rows, classes, lg.stride(0), _p(pr), pr.stride(0), _p(lse), _p(psum),
             _p(g), _p(d), classes, _stream())
        return d, None


def cross_entropy(logits: torch.Tensor, target: torch.Tensor, classes: Optional[int] = None) -> torch.Tensor:
    """F.cross_entropy(logits, target) with mean reduction for 2-D logits; `target` is class indices
    (int64 [rows]) or probabilities (float [rows, classes]). `classes` < logits.shape[1] (index targets): the remaining
    columns are padding -- ignored, and given a zero gradient."""
    if classes is not None and logits.dim() == 2 and target.dtype == torch.int64 and target.dim() == 1:
        return _HLCEFn.apply(logits, target, classes, None)[0]
    if logits.dim() == 3:
        # torch's convention for [N, C, d1] inputs: the CLASS dimension is dim 1 (this is what the reference's
        # IITBehaviorModelPair.loss_fn does to [B, S, V] transformer logits, iit_behavior_model_pair.py:20-30); the mean
        # runs over N * d1 rows
        n, c, d1 = logits.shape
        rows = logits.permute(0, 2, 1).reshape(n * d1, c)
        if target.dtype == torch.int64 and target.shape == (n, d1):
            return _CEIndexFn.apply(rows, target.reshape(n * d1))
        if target.is_floating_point() and target.shape == logits.shape:
            return _CEProbFn.apply(rows, target.permute(0, 2, 1).reshape(n * d1, c))
        raise ValueError(f"cross_entropy: unsupported target {target.dtype} {tuple(target.shape)} for logits {tuple(logits.shape)}")
    if target.dtype == torch.int64 and target.dim() == 1:
        return _CEIndexFn.apply(logits, target)
    if target.is_floating_point() and target.shape == logits.shape:
        return _CEProbFn.apply(logits, target)
    raise ValueError(f"cross_entropy: unsupported target {target.dtype} {tuple(target.shape)} for logits "
                     f"{tuple(logits.shape)}")


# =================================================================================================
# (d) IIA
# =================================================================================================
def iia_count(logits: torch.Tensor, hl: Optional[torch.Tensor], changed: Optional[torch.Tensor] = None,
              counts: Optional[torch.Tensor] = None, want_top1: bool = False):
    """Accumulates (hits, denominator) into `counts` (int64[2], created if None); optionally returns
    top-1 indices. No host sync."""
    _require_cuda(logits, hl, changed, counts)
    logits = logits.detach()
    if logits.dim() > 2:        # [..., classes] (e.g. [B, S, V] transformer logits): one row per leading index
        logits = logits.reshape(-1, logits.shape[-1])
        hl = hl.reshape(-1) if hl is not None else None
        changed = changed.reshape(-1) if changed is not None else None
    lg, rs = _rows_view(logits)
    rows, classes = lg.shape
    if counts is None and hl is not None:
        counts = torch.zeros(2, dtype=torch.int64, device=lg.device)
    top1 = torch.empty(rows, dtype=torch.int64, device=lg.device) if want_top1 else None
    hl_c = hl.contiguous() if hl is not None else None
    if hl_c is not None and (hl_c.dtype != torch.int64 or hl_c.numel() != rows):
        raise ValueError("iia_count: hl must be int64 with one entry per row")
    ch = None
    if changed is not None:
        ch = changed.to(torch.uint8).contiguous() if changed.dtype != torch.uint8 else changed.contiguous()
    call("iit_iia_count", _p(lg), _dt(lg), rows, classes, rs, _p(hl_c), _p(ch), _p(top1), _p(counts), _stream())
    return counts, top1


def argmax_rows(logits: torch.Tensor) -> torch.Tensor:
    return iia_count(logits, None, want_top1=True)[1]


# =================================================================================================
# (b) ResNet layers
# =================================================================================================
def _conv_desc(x, k, r, s, stride, pad) -> Conv:
    n, c, h, w = x.shape
    d = Conv()
    d.n, d.h, d.w, d.c, d.k, d.r, d.s, d.stride, d.pad = n, h, w, c, k, r, s, stride, pad
    d.p = (h + 2 * pad - r) // stride + 1
    d.q = (w + 2 * pad - s) // stride + 1
    return d


def filter_relayout(weight: torch.Tensor, to_crsk: bool, dtype) -> torch.Tensor:
    """fp32 master filter [K,C,R,S] (any strides) -> dense KRSC or CRSK buffer in `dtype`."""
    _require_cuda(weight)
    k, c, r, s = weight.shape
    out = torch.empty(k * c * r * s, dtype=dtype, device=weight.device)
    strides = (C.c_int64 * 4)(*weight.stride())
    call("iit_filter_relayout", _p(weight), strides, _p(out), k, c, r, s, int(to_crsk), _DT[dtype], _stream())
    return out


def filter_relayout_batched(table_dev: torch.Tensor, entries: int, max_elements: int):
    """Every (filter, layout) pair described by the device-resident iit_relayout_t table, in one launch."""
    call("iit_filter_relayout_batched", _p(table_dev), int(entries), int(max_elements), _stream())


def direct_grad(p: torch.Tensor) -> Optional[torch.Tensor]:
    """The gradient buffer of parameter `p` when backward kernels may add into it in place: the view of the optimizer's
    flat gradient buffer that FlatAdam installed as `p.grad` (and marked), still in place. A backward that accumulates there
    returns None for that input, so autograd launches no AccumulateGrad addition. None in every other case."""
    g = p.grad
    if g is None or getattr(p, "_iit_direct_grad", None) is not g or g.dtype != torch.float32 or (g.data_ptr() & 15):
        return None
    return g


USE_GRAD_SLOTS = True     # tests flip it to compare with autograd's own addition


class GradSlot:
    """Meeting point for the two gradients of a tensor that feeds two consumers inside one residual block (conv1 and the
    skip path). Whichever backward runs first writes its gradient as usual, returns it to autograd and leaves the buffer
    here; the one that runs second ADDS its result onto that buffer inside its own kernel (iit_conv2d_dgrad_tc_accumulate)
    and returns None -- so autograd, which holds the first tensor by reference until both are done, never launches an
    element-wise addition. A consumer that cannot accumulate (another dtype / kernel family) ignores the slot and returns
    its gradient normally; autograd then adds, as before."""
    __slots__ = ("_buf", "_task")

    def __init__(self):
        self._buf, self._task = None, -1

    @staticmethod
    def _current_task() -> int:
        return int(torch._C._current_graph_task_id())      # identifies the backward pass that is running (-1: none)

    def put(self, t: torch.Tensor):
        self._buf, self._task = t, self._current_task()

    def take(self):
        """The gradient the other consumer left during THIS backward pass (a leftover of an earlier, incomplete pass is
        dropped), or None."""
        buf, task = self._buf, self._task
        self._buf, self._task = None, -1
        return buf if (buf is not None and task == self._current_task() and task != -1) else None


class _ConvFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, weight, w_krsc, w_crsk_fn, stride, pad, bn_sums, wgrad_buf, slot=None, patch=None):
        _require_cuda(x, weight)
        x = _cl(x)
        k, c, r, s = weight.shape
        d = _conv_desc(x, k, r, s, stride, pad)
        y = empty_cl(d.n, k, d.p, d.q, x.dtype, x.device)
        s0 = s1 = None
        if bn_sums is not None:
            s0, s1 = bn_sums[0], bn_sums[1]
        tc = USE_TENSOR_CORES and bool(_lib.lib.iit_conv2d_tc_supported(C.byref(d), _dt(x)))
        if patch is not None:   # (cached source activation, box6): the interchange runs in this convolution's epilogue
            src, box6 = patch
            if not tc or src.shape != y.shape or src.dtype != y.dtype or not _is_cl(src):
                raise RuntimeError("conv2d: a fused patch needs the tensor-core path and a source activation laid out like the output")
            call("iit_conv2d_fprop_tc_patch", _p(x), _p(w_krsc), _p(y), C.byref(d), _p(s0), _p(s1), _p(src),
                 (C.c_int32 * 6)(*box6), _stream())
        elif tc:  # tcgen05 + TMA implicit GEMM; BN statistics (if wanted) by the statistics kernel
            call("iit_conv2d_fprop_tc", _p(x), _p(w_krsc), _p(y), C.byref(d), _p(s0), _p(s1), _stream())
        else:
            call("iit_conv2d_fprop", _p(x), _p(w_krsc), _p(y), C.byref(d), _dt(x), _p(s0), _p(s1), _stream())
        ctx.save_for_backward(x, weight)
        ctx.d, ctx.w_crsk_fn, ctx.tc = d, w_crsk_fn, tc
        ctx.slot = slot if tc else None
        # physical KRSC (channels-last filter) gradient buffer the weight-gradient reduction may add into directly
        ctx.wgrad_buf = wgrad_buf if (tc and wgrad_buf is not None and wgrad_buf.shape == weight.shape
                                      and wgrad_buf.permute(0, 2, 3, 1).is_contiguous()) else None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, weight = ctx.saved_tensors
        d = ctx.d
        dy = _cl(dy)
        dx = dw = None
        if ctx.needs_input_grad[0]:
            slot = ctx.slot if USE_GRAD_SLOTS else None
            first = slot.take() if slot is not None else None
            if first is not None and first.shape == x.shape and first.dtype == x.dtype and _is_cl(first):
                # the other consumer's gradient is already there: add this one onto it, hand autograd nothing
                call("iit_conv2d_dgrad_tc_accumulate", _p(dy), _p(ctx.w_crsk_fn()), _p(first), C.byref(d), _stream())
            else:
                dx = empty_cl(d.n, d.c, d.h, d.w, x.dtype, x.device)
                if ctx.tc:
                    call("iit_conv2d_dgrad_tc", _p(dy), _p(ctx.w_crsk_fn()), _p(dx), C.byref(d), _stream())
                    if slot is not None:
                        slot.put(dx)
                else:
                    call("iit_conv2d_dgrad", _p(dy), _p(ctx.w_crsk_fn()), _p(dx), C.byref(d), _dt(x), _stream())
        if ctx.needs_input_grad[1] and ctx.wgrad_buf is not None:
            need = int(_lib.lib.iit_conv2d_wgrad_tc_scratch_bytes(C.byref(d)))
            scratch = torch.empty(need, dtype=torch.uint8, device=x.device)
            call("iit_conv2d_wgrad_tc_accumulate", _p(dy), _p(x), _p(ctx.wgrad_buf), C.byref(d), _p(scratch), need, _stream())
        elif ctx.needs_input_grad[1]:
            dwb = torch.empty((d.k, d.r, d.s, d.c), dtype=torch.float32, device=x.device)
            if ctx.tc:
                need = int(_lib.lib.iit_conv2d_wgrad_tc_scratch_bytes(C.byref(d)))
                scratch = torch.empty(need, dtype=torch.uint8, device=x.device)
                call("iit_conv2d_wgrad_tc", _p(dy), _p(x), _p(dwb), C.byref(d), _p(scratch), need, _stream())
            else:
                need = int(_lib.lib.iit_conv2d_wgrad_scratch_bytes(C.byref(d)))
                scratch = torch.empty(need, dtype=torch.uint8, device=x.device) if need else None
                call("iit_conv2d_wgrad", _p(dy), _p(x), _p(dwb), C.byref(d), _dt(x), _p(scratch), need, _stream())
            dw = dwb.permute(0, 3, 1, 2)  # [K,C,R,S] view with channels-last strides, like the parameter
        return dx, dw, None, None, None, None, None, None, None, None


FOLD_EVAL_BN = True     # evaluation passes run conv + BatchNorm (+ residual) (+ ReLU) as one kernel (tests flip it to compare)


def conv_affine_ok(x, weight, stride, pad) -> bool:
    """Whether conv2d_affine can take this convolution (the tensor-core path on bf16 activations, outside autograd)."""
    if not (FOLD_EVAL_BN and USE_TENSOR_CORES and x.is_cuda and x.dtype == torch.bfloat16 and not torch.is_grad_enabled()):
        return False
    k, c, r, s = weight.shape
    return bool(_lib.lib.iit_conv2d_tc_supported(C.byref(_conv_desc(_cl(x), k, r, s, stride, pad)), _dt(x)))


def conv2d_affine(x, weight_shape, w_krsc, stride, pad, scale, shift, residual=None, relu=False):
    """y = act(conv(x, w) * scale[k] + shift[k] (+ residual)) in one kernel (iit_conv2d_fprop_tc_affine): an evaluation-mode
    BatchNorm, the block's residual addition and its ReLU folded into the convolution's epilogue. No autograd: evaluation
    passes only (the causality sweep, run_eval_step)."""
    _require_cuda(x, scale, shift, residual)
    x = _cl(x)
    k, c, r, s = weight_shape
    d = _conv_desc(x, k, r, s, stride, pad)
    y = empty_cl(d.n, k, d.p, d.q, x.dtype, x.device)
    if residual is not None:
        residual = _cl(residual)
        if residual.shape != y.shape or residual.dtype != y.dtype:
            raise RuntimeError("conv2d_affine: the residual must have the output's shape and dtype")
    call("iit_conv2d_fprop_tc_affine", _p(x), _p(w_krsc), _p(y), C.byref(d), _p(scale), _p(shift), _p(residual), int(bool(relu)), _stream())
    return y


def bn_eval_affine(gamma, beta, running_mean, running_var, eps):
    """(scale, shift) of an evaluation-mode BatchNorm as one per-channel affine map."""
    c = gamma.numel()
    scale = torch.empty(c, dtype=torch.float32, device=gamma.device)
    shift = torch.empty(c, dtype=torch.float32, device=gamma.device)
    call("iit_bn_eval_affine", _p(gamma), _p(beta), _p(running_mean), _p(running_var), c, float(eps), _p(scale), _p(shift), _stream())
    return scale, shift


def conv2d(x, weight, w_krsc, w_crsk_fn, stride, pad, bn_sums=None, wgrad_buf=None, slot=None, patch=None):
    """`wgrad_buf`: see direct_grad -- the weight gradient is added into it and autograd receives None.
    `slot`: see GradSlot -- shared with the other consumer of `x` in the same residual block.
    `patch`: (src, (p0, p1, q0, q1, c0, c1)) -- output rows / columns / channels inside the box take src's values, in the
    kernel's epilogue (iit_conv2d_fprop_tc_patch). The caller must see to it that the gradient reaching this convolution is
    zero inside the box (bn_act(..., zero_box=...) does it in the BatchNorm backward's dx store)."""
    return _ConvFn.apply(x, weight, w_krsc, w_crsk_fn, stride, pad, bn_sums, wgrad_buf, slot, patch)


def conv_out_shape(x, weight, stride, pad):
    n, _, h, w = x.shape
    k, _, r, s_ = weight.shape
    return (n, k, (h + 2 * pad - r) // stride + 1, (w + 2 * pad - s_) // stride + 1)


def conv_patch_fusable(x, weight, stride, pad) -> bool:
    if not (USE_TENSOR_CORES and x.is_cuda and x.dtype == torch.bfloat16):
        return False
    d = _conv_desc(x, weight.shape[0], weight.shape[2], weight.shape[3], stride, pad)
    return bool(_lib.lib.iit_conv2d_patch_fusable(C.byref(d)))


STEM_K = 192


def stem_im2col(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """[N,3,H,W] image (fp32 or bf16, any strides) -> bf16 [N,192,P,Q] channels-last im2col tensor of the stem."""
    _require_cuda(x)
    n, c, h, w = x.shape
    if c != 3:
        raise ValueError("stem_im2col expects 3 input channels")
    p, q = (h + 6 - 7) // 2 + 1, (w + 6 - 7) // 2 + 1
    col = out if out is not None else empty_cl(n, STEM_K, p, q, torch.bfloat16, x.device)
    strides = (C.c_int64 * 4)(*x.stride())
    call("iit_stem_im2col", _p(x), _dt(x), strides, _p(col), n, h, w, _stream())
    return col


# Operands derived from the image batch alone (not from the weights) -- the stem's space-to-depth image for the forward,
# its im2col matrix for the weight gradient -- are built once per batch tensor and kept while that tensor object is alive
# and unmodified (`_version`): the five LL forwards of one train step see only two distinct batches (base, source).
# Under CUDA-graph capture a registered tensor is a static buffer: the captured kernels read the derived buffer, which
# the caller refreshes eagerly (`refresh_batch_operands`, outside capture) before replaying.
_batch_cache: dict = {}


def _batch_cached(x: torch.Tensor, kind: str, build):
    key = (id(x), kind)
    ent = _batch_cache.get(key)
    if ent is not None and ent[0]() is x and ent[2].shape[0] == x.shape[0]:
        if capturing() or ent[1] == x._version:
            return ent[2]
        build(x, ent[2])
        ent[1] = x._version
        return ent[2]
    if capturing():
        return build(x, None)
    buf = build(x, None)
    _batch_cache[key] = [weakref.ref(x, lambda _r, k=key: _batch_cache.pop(k, None)), x._version, buf]
    return buf


def refresh_batch_operands(x: torch.Tensor):
    """Rebuild every operand already derived from `x` (after its contents changed in place)."""
    for kind, build in (("s2d", stem_s2d), ("col", stem_im2col)):
        if (id(x), kind) in _batch_cache:
            _batch_cached(x, kind, build)


def stem_s2d(x: torch.Tensor, out: Optional[torch.Tensor] = None) -> torch.Tensor:
    """[N,3,H,W] image (fp32 or bf16, any strides; H, W even) -> bf16 [N,(H+6)/2,(W+6)/2,16]: zero-padded by 3 and 2x2
    space-to-depth, channel (a*2+b)*3+c <- padded[c, 2y+a, 2x+b]."""
    _require_cuda(x)
    n, c, h, w = x.shape
    if c != 3 or h % 2 or w % 2:
        raise ValueError("stem_s2d expects 3 input channels and even height / width")
    buf = out if out is not None else torch.empty((n, (h + 6) // 2, (w + 6) // 2, 16), dtype=torch.bfloat16, device=x.device)
    strides = (C.c_int64 * 4)(*x.stride())
    call("iit_stem_s2d", _p(x), _dt(x), strides, _p(buf), n, h, w, _stream())
    return buf


def stem_filter_s2d(weight: torch.Tensor) -> torch.Tensor:
    """fp32 [64,3,7,7] filter (any strides) -> bf16 [64,4,4,16] in the space-to-depth grouping."""
    _require_cuda(weight)
    if tuple(weight.shape) != (64, 3, 7, 7) or weight.dtype != torch.float32:
        raise ValueError("stem_filter_s2d expects the fp32 [64,3,7,7] stem filter")
    w2 = torch.empty((64, 4, 4, 16), dtype=torch.bfloat16, device=weight.device)
    strides = (C.c_int64 * 4)(*weight.stride())
    call("iit_stem_filter_s2d", _p(weight), strides, _p(w2), _stream())
    return w2


def stem_cols(x: torch.Tensor) -> torch.Tensor:
    return _batch_cached(x, "col", stem_im2col)


def stem_s2d_cached(x: torch.Tensor) -> torch.Tensor:
    return _batch_cached(x, "s2d", stem_s2d)


class _StemConvFn(torch.autograd.Function):
    """7x7/2 stem on the tensor cores. Forward: the 4x4 stride-1 convolution over the space-to-depth image on the
    shifted-window kernel (even image sizes with 64 filters), else im2col + a 1x1 convolution with 192 channels.
    Weight gradient: the shifted-window kernel on the same space-to-depth image (s2d width <= 32), else im2col (built
    once per batch) + the 1x1 wgrad kernel."""

    @staticmethod
    def forward(ctx, x, weight, w_prepared, bn_sums):
        _require_cuda(x, weight)
        n, _, h, w = x.shape
        k = weight.shape[0]
        s0 = s1 = None
        if bn_sums is not None:
            s0, s1 = bn_sums[0], bn_sums[1]
        if w_prepared.dim() == 4:     # [64,4,4,16]: space-to-depth form
            s2d = stem_s2d_cached(x)
            y = empty_cl(n, k, h // 2, w // 2, torch.bfloat16, x.device)
            call("iit_stem_conv_s2d", _p(s2d), _p(w_prepared), _p(y), n, s2d.shape[1], s2d.shape[2], _p(s0), _p(s1), _stream())
        else:                          # [K,192]: im2col form
            col = stem_cols(x)
            _, _, p, q = col.shape
            d = _conv_desc(col, k, 1, 1, 1, 0)
            y = empty_cl(n, k, p, q, torch.bfloat16, x.device)
            call("iit_conv2d_fprop_tc", _p(col), _p(w_prepared), _p(y), C.byref(d), _p(s0), _p(s1), _stream())
        ctx.x = x                      # not a saved tensor: the batch is an input, never produced by autograd
        ctx.save_for_backward(weight)
        return y

    @staticmethod
    def backward(ctx, dy):
        (weight,) = ctx.saved_tensors
        dw = None
        if ctx.needs_input_grad[1]:
            x = ctx.x
            n, _, h, w = x.shape
            dy = _cl(dy)
            sh, sw = (h + 6) // 2, (w + 6) // 2
            need = int(_lib.lib.iit_stem_wgrad_s2d_scratch_bytes(n, sh, sw)) if (h % 2 == 0 and w % 2 == 0 and weight.shape[0] == 64) else 0
            if need:   # shifted-window kernel on the space-to-depth image (already built for the forward)
                s2d = stem_s2d_cached(x)
                dw = torch.empty_like(weight, memory_format=torch.preserve_format)
                scratch = torch.empty(need, dtype=torch.uint8, device=x.device)
                strides = (C.c_int64 * 4)(*dw.stride())
                call("iit_stem_wgrad_s2d", _p(dy), _p(s2d), _p(dw), strides, n, sh, sw, _p(scratch), need, _stream())
            else:      # im2col (built once per batch) + the 1x1 wgrad kernel with 192 channels
                col = stem_cols(x)
                k = weight.shape[0]
                d = _conv_desc(col, k, 1, 1, 1, 0)
                dw192 = torch.empty((d.k, STEM_K), dtype=torch.float32, device=col.device)
                need = int(_lib.lib.iit_conv2d_wgrad_tc_scratch_bytes(C.byref(d)))
                scratch = torch.empty(need, dtype=torch.uint8, device=col.device)
                call("iit_conv2d_wgrad_tc", _p(dy), _p(col), _p(dw192), C.byref(d), _p(scratch), need, _stream())
                dw = dw192[:, :147].reshape(d.k, 7, 7, 3).permute(0, 3, 1, 2)
        return None, dw, None, None


def stem_conv(x, weight, w192, bn_sums=None):
    return _StemConvFn.apply(x, weight, w192, bn_sums)


def new_bn_sums(c: int, device) -> torch.Tensor:
    return torch.zeros((2, c), dtype=torch.float64, device=device)


BN_RELU_MASK = True     # BatchNorm+ReLU forward keeps a bit mask of the positive outputs for its backward (tests flip it)


class _BNActFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, residual, running_mean, running_var, training, momentum, eps, relu, sums, acc, tracked,
                bacc, pgrads, slot=None, zero_box=None):
        _require_cuda(x, gamma, beta, residual)
        x = _cl(x)
        n, c, h, w = x.shape
        rows = n * h * w
        st = _stream()
        if residual is not None:
            residual = _cl(residual)
        if training:
            clear = 0
            if sums is None:
                # `acc`: the layer's persistent accumulator (zero between uses: the finalize kernel clears it)
                sums, clear = (acc, 1) if acc is not None else (new_bn_sums(c, x.device), 0)
                call("iit_bn_stats", _p(x), _dt(x), rows, c, _p(sums[0]), _p(sums[1]), st)
            elif acc is not None and sums.data_ptr() == acc.data_ptr():
                clear = 1
            mean = torch.empty(c, dtype=torch.float32, device=x.device)
            invstd = torch.empty(c, dtype=torch.float32, device=x.device)
            call("iit_bn_finalize", _p(sums[0]), _p(sums[1]), rows, c, float(eps), float(momentum), _p(mean), _p(invstd),
                 _p(running_mean), _p(running_var), _p(tracked), clear, st)
        else:
            mean = running_mean
            invstd = torch.empty(c, dtype=torch.float32, device=x.device)
            call("iit_bn_eval_invstd", _p(running_var), c, float(eps), _p(invstd), st)
        y = torch.empty_like(x)
        # ReLU mask for the backward, one bit per element (it then reads 1/16 of what y would cost): only when one follows
        mask = None
        if BN_RELU_MASK and relu and any(ctx.needs_input_grad[:4]):
            mask = torch.empty(rows * c // (8 if x.dtype == torch.bfloat16 else 4), dtype=torch.uint8, device=x.device)
        call("iit_bn_act_fwd", _p(x), _p(residual), _p(y), _dt(x), rows, c, _p(mean), _p(invstd), _p(gamma), _p(beta),
             int(relu), _p(mask), st)
        ctx.save_for_backward(x, y if relu else None, gamma, mean, invstd, mask)
        ctx.relu, ctx.training, ctx.has_res = relu, training, residual is not None
        ctx.bacc, ctx.pgrads = bacc, pgrads
        ctx.slot = slot if residual is not None else None
        ctx.zero_box = zero_box
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y, gamma, mean, invstd, mask = ctx.saved_tensors
        dy = _cl(dy)
        n, c, h, w = x.shape
        rows = n * h * w
        st = _stream()
        # `bacc`: the layer's persistent fp64 [2, C] accumulator (zero between uses; iit_bn_param_grads clears it)
        red = ctx.bacc if ctx.bacc is not None else torch.zeros((2, c), dtype=torch.float64, device=x.device)
        try:
            call("iit_bn_act_bwd_reduce", _p(dy), _p(y), _p(x), _dt(x), rows, c, _p(mean), _p(invstd), int(ctx.relu),
                 _p(red[0]), _p(red[1]), _p(mask), st)
            dx = torch.empty_like(x)
            dres = torch.empty_like(x) if ctx.has_res else None
            # eval mode: statistics are constants, so the two correction terms vanish
            corr = red if ctx.training else torch.zeros_like(red)
            if ctx.zero_box is not None:    # x was patched inside its convolution: zero gradient inside the box, written here
                call("iit_bn_act_bwd_apply_box", _p(dy), _p(y), _p(x), _p(dx), _p(dres), _dt(x), rows, c, _p(mean), _p(invstd),
                     _p(gamma), _p(corr[0]), _p(corr[1]), int(ctx.relu), _p(mask), h, w, (C.c_int32 * 6)(*ctx.zero_box), st)
            else:
                call("iit_bn_act_bwd_apply", _p(dy), _p(y), _p(x), _p(dx), _p(dres), _dt(x), rows, c, _p(mean), _p(invstd),
                     _p(gamma), _p(corr[0]), _p(corr[1]), int(ctx.relu), _p(mask), st)
            dgamma, dbeta = _bn_param_grads(red, c, ctx.pgrads, ctx.bacc is not None, st)
        except Exception:
            if ctx.bacc is not None:
                ctx.bacc.zero_()
            raise
        if ctx.slot is not None and dres is not None and USE_GRAD_SLOTS:
            ctx.slot.put(dres)       # the block input's other consumer (conv1) adds its gradient onto this (GradSlot)
        return dx, dgamma, dbeta, dres, None, None, None, None, None, None, None, None, None, None, None, None, None


def _bn_param_grads(red, c, pgrads, clear, st):
    """fp64 sums -> the fp32 gradients of (gamma, beta) in one launch: added into the optimizer's buffers (`pgrads`, see
    direct_grad; autograd then receives None) or written to fresh tensors."""
    if pgrads is not None:
        call("iit_bn_param_grads", _p(red), c, _p(pgrads[0]), _p(pgrads[1]), 1, int(clear), st)
        return None, None
    out = torch.empty((2, c), dtype=torch.float32, device=red.device)
    call("iit_bn_param_grads", _p(red), c, _p(out[0]), _p(out[1]), 0, int(clear), st)
    return out[0], out[1]


def bn_act(x, gamma, beta, residual, running_mean, running_var, training, momentum, eps, relu, sums=None, acc=None,
           tracked=None, bacc=None, pgrads=None, slot=None, zero_box=None):
    """`sums`: statistics already accumulated by the producing convolution's epilogue; `acc`: the layer's persistent
    fp64 accumulator [2, C] (kept zero between uses; `sums` may be it); `tracked`: num_batches_tracked, incremented by the
    finalize kernel; `bacc`: the same kind of accumulator for the backward's sums; `pgrads`: (gamma, beta) gradient buffers
    to add into (direct_grad)."""
    return _BNActFn.apply(x, gamma, beta, residual, running_mean, running_var, training, momentum, eps, relu, sums, acc, tracked,
                          bacc, pgrads, slot, zero_box)


class _BNReluPoolFn(torch.autograd.Function):
    """bn (train or eval statistics) -> relu -> maxpool 3x3/2 of the stem conv output as one kernel. Stored for the
    backward: the window index and the winners' raw values (pooled size); dx recomputes the activation from the conv output."""

    @staticmethod
    def forward(ctx, x, gamma, beta, running_mean, running_var, training, momentum, eps, sums, acc, tracked, bacc, pgrads):
        _require_cuda(x, gamma, beta)
        x = _cl(x)
        n, c, h, w = x.shape
        rows = n * h * w
        st = _stream()
        if training:
            clear = 0
            if sums is None:
                sums, clear = (acc, 1) if acc is not None else (new_bn_sums(c, x.device), 0)
                call("iit_bn_stats", _p(x), _dt(x), rows, c, _p(sums[0]), _p(sums[1]), st)
            elif acc is not None and sums.data_ptr() == acc.data_ptr():
                clear = 1
            mean = torch.empty(c, dtype=torch.float32, device=x.device)
            invstd = torch.empty(c, dtype=torch.float32, device=x.device)
            call("iit_bn_finalize", _p(sums[0]), _p(sums[1]), rows, c, float(eps), float(momentum), _p(mean), _p(invstd),
                 _p(running_mean), _p(running_var), _p(tracked), clear, st)
        else:
            mean = running_mean
            invstd = torch.empty(c, dtype=torch.float32, device=x.device)
            call("iit_bn_eval_invstd", _p(running_var), c, float(eps), _p(invstd), st)
        p, q = (h + 2 - 3) // 2 + 1, (w + 2 - 3) // 2 + 1
        y = empty_cl(n, c, p, q, x.dtype, x.device)
        idx = torch.empty((n, p, q, c), dtype=torch.int8, device=x.device)
        # the winners' raw values, kept only when a backward will follow: its per-channel sums then run over pooled tensors
        xwin = torch.empty_like(y) if any(ctx.needs_input_grad[:3]) else None
        call("iit_bn_relu_maxpool_fwd", _p(x), _p(y), _p(idx), _p(xwin), _dt(x), n, h, w, c, _p(mean), _p(invstd), _p(gamma),
             _p(beta), st)
        ctx.save_for_backward(x, idx, gamma, beta, mean, invstd, y, xwin)
        ctx.training = training
        ctx.bacc, ctx.pgrads = bacc, pgrads
        return y

    @staticmethod
    def backward(ctx, dy):
        x, idx, gamma, beta, mean, invstd, y, xwin = ctx.saved_tensors
        dy = _cl(dy)
        n, c, h, w = x.shape
        if not ctx.training:
            raise NotImplementedError("fused bn+relu+maxpool backward is for batch statistics (train mode)")
        red = ctx.bacc if ctx.bacc is not None else torch.zeros((2, c), dtype=torch.float64, device=x.device)
        dx = torch.empty_like(x)
        st = _stream()
        try:
            call("iit_bn_relu_maxpool_bwd", _p(dy), _p(idx), _p(x), _p(y), _p(xwin), _p(dx), _dt(x), n, h, w, c, _p(mean),
                 _p(invstd), _p(gamma), _p(beta), _p(red[0]), _p(red[1]), st)
            dgamma, dbeta = _bn_param_grads(red, c, ctx.pgrads, ctx.bacc is not None, st)
        except Exception:
            if ctx.bacc is not None:
                ctx.bacc.zero_()
            raise
        return dx, dgamma, dbeta, None, None, None, None, None, None, None, None, None, None


def bn_relu_maxpool(x, gamma, beta, running_mean, running_var, training, momentum, eps, sums=None, acc=None, tracked=None,
                    bacc=None, pgrads=None):
    return _BNReluPoolFn.apply(x, gamma, beta, running_mean, running_var, training, momentum, eps, sums, acc, tracked, bacc,
                               pgrads)


class _DivScalarFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, scale):
        _require_cuda(x)
        xc = x if (x.is_contiguous() or _is_cl(x)) else x.contiguous()
        y = torch.empty_like(xc)
        call("iit_div_scalar", _p(xc), _p(y), _dt(xc), xc.numel(), float(scale), _stream())
        ctx.scale = float(scale)
        return y

    @staticmethod
    def backward(ctx, dy):
        dy = dy if (dy.is_contiguous() or _is_cl(dy)) else dy.contiguous()
        dx = torch.empty_like(dy)
        call("iit_div_scalar", _p(dy), _p(dx), _dt(dy), dy.numel(), ctx.scale, _stream())
        return dx, None


def div_scalar(x, scale: float):
    """x / scale with autograd (the stop-gradient pair's forward hook, stop_grad_pair.py:37-44)."""
    return _DivScalarFn.apply(x, scale)


class _ReluFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        _require_cuda(x)
        xc = x if (x.is_contiguous() or _is_cl(x)) else x.contiguous()
        y = torch.empty_like(xc)
        call("iit_relu_fwd", _p(xc), _p(y), _dt(xc), xc.numel(), _stream())
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        dy = dy.contiguous(memory_format=torch.channels_last) if _is_cl(y) else dy.contiguous()
        dx = torch.empty_like(y)
        call("iit_relu_bwd", _p(dy), _p(y), _p(dx), _dt(y), y.numel(), _stream())
        return dx


def relu(x):
    return _ReluFn.apply(x)


class _AddReluFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b, do_relu):
        _require_cuda(a, b)
        a, b = _cl(a), _cl(b)
        y = torch.empty_like(a)
        call("iit_add_relu_fwd", _p(a), _p(b), _p(y), _dt(a), a.numel(), int(do_relu), _stream())
        ctx.do_relu = do_relu
        ctx.save_for_backward(y if do_relu else None)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        if not ctx.do_relu:
            return dy, dy, None
        dy = _cl(dy)
        dx = torch.empty_like(y)
        call("iit_relu_bwd", _p(dy), _p(y), _p(dx), _dt(y), y.numel(), _stream())
        return dx, dx, None


def add_relu(a, b, do_relu=True):
    return _AddReluFn.apply(a, b, do_relu)


class _MaxPoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        _require_cuda(x)
        x = _cl(x)
        n, c, h, w = x.shape
        p, q = (h + 2 - 3) // 2 + 1, (w + 2 - 3) // 2 + 1
        y = empty_cl(n, c, p, q, x.dtype, x.device)
        idx = torch.empty((n, p, q, c), dtype=torch.int8, device=x.device)
        call("iit_maxpool3x3s2_fwd", _p(x), _p(y), _p(idx), _dt(x), n, h, w, c, _stream())
        ctx.save_for_backward(idx)
        ctx.shape = (n, c, h, w)
        return y

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        n, c, h, w = ctx.shape
        dy = _cl(dy)
        dx = empty_cl(n, c, h, w, dy.dtype, dy.device)
        call("iit_maxpool3x3s2_bwd", _p(dy), _p(idx), _p(dx), _dt(dy), n, h, w, c, _stream())
        return dx


def maxpool3x3s2(x):
    return _MaxPoolFn.apply(x)


class _AvgPoolFn(torch.autograd.Function):
    """AdaptiveAvgPool2d(1): [N,C,H,W] -> fp32 [N,C,1,1]."""

    @staticmethod
    def forward(ctx, x):
        _require_cuda(x)
        x = _cl(x)
        n, c, h, w = x.shape
        y = torch.empty((n, c), dtype=torch.float32, device=x.device)
        call("iit_avgpool_fwd", _p(x), _p(y), _dt(x), n, h * w, c, _stream())
        ctx.shape, ctx.dtype = (n, c, h, w), x.dtype
        return y.view(n, c, 1, 1)

    @staticmethod
    def backward(ctx, dy):
        n, c, h, w = ctx.shape
        dy = dy.reshape(n, c).float().contiguous()
        dx = empty_cl(n, c, h, w, ctx.dtype, dy.device)
        call("iit_avgpool_bwd", _p(dy), _p(dx), _DT[ctx.dtype], n, h * w, c, _stream())
        return dx


def avgpool(x):
    return _AvgPoolFn.apply(x)


class _LinearFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, w, b):
        _require_cuda(x, w, b)
        x = x.float().contiguous()
        wc = w.contiguous()
        n, i = x.shape
        o = wc.shape[0]
        y = torch.empty((n, o), dtype=torch.float32, device=x.device)
        call("iit_linear_fwd", _p(x), _p(wc), _p(b), _p(y), n, i, o, _stream())
        ctx.save_for_backward(x, wc)
        ctx.has_bias = b is not None
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        dy = dy.float().contiguous()
        n, i = x.shape
        o = w.shape[0]
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        dw = torch.empty_like(w)
        db = torch.empty(o, dtype=torch.float32, device=x.device) if ctx.has_bias else None
        call("iit_linear_bwd", _p(x), _p(w), _p(dy), _p(dx), _p(dw), _p(db), n, i, o, _stream())
        return dx, dw, db


def linear(x, w, b):
    return _LinearFn.apply(x, w, b)


# =================================================================================================
# (b) transformer layers
# =================================================================================================
def _as_filter4(w_in_out: torch.Tensor) -> torch.Tensor:
    """[in, out] weight (transformer_lens layout) viewed as a [K=out, C=in, 1, 1] filter (no copy)."""
    return w_in_out.t().unsqueeze(-1).unsqueeze(-1)


class _LinearTokensFn(torch.autograd.Function):
    """y[T, out] = x[T, in] @ W[in, out]: a 1x1 convolution over T 'pixels' on the convolution GEMM kernels
    (tcgen05 when bf16 and both widths are multiples of 64, the exact-fp32 CUDA-core GEMM otherwise)."""

    @staticmethod
    def forward(ctx, x, W):
        _require_cuda(x, W)
        x = x.contiguous()
        T, cin = x.shape
        cout = W.shape[1]
        d = Conv()
        d.n, d.h, d.w, d.c, d.k, d.r, d.s, d.stride, d.pad, d.p, d.q = T, 1, 1, cin, cout, 1, 1, 1, 0, 1, 1
        tc = USE_TENSOR_CORES and bool(_lib.lib.iit_conv2d_tc_supported(C.byref(d), _dt(x)))
        w_oi = filter_relayout(_as_filter4(W.detach()), False, x.dtype)      # dense [out][in]
        y = torch.empty((T, cout), dtype=x.dtype, device=x.device)
        if tc:
            call("iit_conv2d_fprop_tc", _p(x), _p(w_oi), _p(y), C.byref(d), None, None, _stream())
        else:
            call("iit_conv2d_fprop", _p(x), _p(w_oi), _p(y), C.byref(d), _dt(x), None, None, _stream())
        ctx.save_for_backward(x, W)
        ctx.d, ctx.tc = d, tc
        return y

    @staticmethod
    def backward(ctx, dy):
        x, W = ctx.saved_tensors
        d = ctx.d
        dy = dy.contiguous()
        dx = dW = None
        if ctx.needs_input_grad[0]:
            # dgrad wants the filter as [in][out]: W itself when it is fp32 and dense
            if x.dtype == torch.float32 and W.is_contiguous():
                w_io = W.detach()
            else:
                w_io = filter_relayout(_as_filter4(W.detach()), True, x.dtype)
            dx = torch.empty_like(x)
            if ctx.tc:
                call("iit_conv2d_dgrad_tc", _p(dy), _p(w_io), _p(dx), C.byref(d), _stream())
            else:
                call("iit_conv2d_dgrad", _p(dy), _p(w_io), _p(dx), C.byref(d), _dt(x), _stream())
        if ctx.needs_input_grad[1]:
            dwb = torch.empty((d.k, d.c), dtype=torch.float32, device=x.device)   # [out][in]
            if ctx.tc:
                need = int(_lib.lib.iit_conv2d_wgrad_tc_scratch_bytes(C.byref(d)))
                scratch = torch.empty(need, dtype=torch.uint8, device=x.device)
                call("iit_conv2d_wgrad_tc", _p(dy), _p(x), _p(dwb), C.byref(d), _p(scratch), need, _stream())
            else:
                need = int(_lib.lib.iit_conv2d_wgrad_scratch_bytes(C.byref(d)))
                scratch = torch.empty(need, dtype=torch.uint8, device=x.device) if need else None
                call("iit_conv2d_wgrad", _p(dy), _p(x), _p(dwb), C.byref(d), _dt(x), _p(scratch), need, _stream())
            dW = dwb.t()                                                       # [in][out] view
        return dx, dW


def linear_tokens(x2d: torch.Tensor, W_in_out: torch.Tensor) -> torch.Tensor:
    return _LinearTokensFn.apply(x2d, W_in_out)


def wgrad_direct_maps(specs, params, cache: dict):
    """Device table of iit_wgrad_map_t that lets the weight-gradient GEMM's split-K reduction add straight into the
    optimizer's gradient views of `params` (direct_grad), each laid out as its `spec` (size, stride, offset) of the
    [out_pad][in] matrix says -- or None when any of them may not be written in place. Returned as ("direct", specs,
    table, n, max_numel), the form _LinearPreparedFn / _LinearEpiFn take in place of `specs`. `cache` belongs to the call
    site: the table is rebuilt only when a gradient buffer moved."""
    if not torch.is_grad_enabled():
        return None
    gs = [direct_grad(p) if p.requires_grad else None for p in params]
    if any(g is None or not g.is_contiguous() for g in gs):
        return None
    key = tuple(g.data_ptr() for g in gs)
    ent = cache.get("wgrad_maps")
    if ent is None or ent[0] != key:
        arr = (_lib.WgradMap * len(gs))()
        biggest = 0
        for m, g, (size, stride, offset) in zip(arr, gs, specs):
            if tuple(g.shape) != tuple(size) or len(size) > 3:
                return None
            pad = 3 - len(size)
            m.dst = g.data_ptr()
            for i in range(3):
                m.size[i] = 1 if i < pad else int(size[i - pad])
                m.stride[i] = 0 if i < pad else int(stride[i - pad])
            m.offset = int(offset)
            biggest = max(biggest, g.numel())
        host = torch.frombuffer(bytearray(bytes(arr)), dtype=torch.uint8)
        ent = cache["wgrad_maps"] = (key, host.to(gs[0].device), len(gs), biggest)
    return ("direct", specs, ent[1], ent[2], ent[3])


def _split_direct(specs):
    if isinstance(specs, tuple) and len(specs) == 5 and isinstance(specs[0], str) and specs[0] == "direct":
        return specs[1], specs[2:]
    return specs, None


class _LinearPreparedFn(torch.autograd.Function):
    """y[T, out_pad] = x[T, in] @ W on the tensor-core GEMMs from PREPARED bf16 operands: `w_oi` [out_pad][in] for the
    forward, `w_io` [in][out_pad] for dgrad -- persistent copies the model refreshes in one launch after every weight
    update (HookedTransformerLL.refresh_filter_shadow), instead of one re-layout launch per GEMM and pass. The weight
    gradient [out_pad][in] is handed to each of `params` as the strided view its `spec` (size, stride, offset) names, so
    fused / padded matrices (QKV from three per-head parameters, the padded unembedding) need no cat / pad in the graph."""

    @staticmethod
    def forward(ctx, x, w_oi, w_io, n_in, n_out_pad, specs, *params):
        _require_cuda(x, w_oi, w_io)
        x = x.contiguous()
        T = x.shape[0]
        d = Conv()
        d.n, d.h, d.w, d.c, d.k, d.r, d.s, d.stride, d.pad, d.p, d.q = T, 1, 1, n_in, n_out_pad, 1, 1, 1, 0, 1, 1
        if not (x.dtype == torch.bfloat16 and _lib.lib.iit_conv2d_tc_supported(C.byref(d), _dt(x))):
            raise RuntimeError("linear_prepared: bf16 activations and widths that are multiples of 64 only")
        y = torch.empty((T, n_out_pad), dtype=x.dtype, device=x.device)
        call("iit_conv2d_fprop_tc", _p(x), _p(w_oi), _p(y), C.byref(d), None, None, _stream())
        ctx.save_for_backward(x, w_io)
        ctx.specs, ctx.w_direct = _split_direct(specs)
        ctx.d = d
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w_io = ctx.saved_tensors
        d = ctx.d
        dy = dy.contiguous()
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            call("iit_conv2d_dgrad_tc", _p(dy), _p(w_io), _p(dx), C.byref(d), _stream())
        grads = [None] * len(ctx.specs)
        if any(ctx.needs_input_grad[6:]):
            grads = _wgrad_into(dy, x, d, ctx.specs, ctx.w_direct)
        return (dx, None, None, None, None, None, *grads)


def _wgrad_into(g, x, d, specs, w_direct):
    """The [out_pad][in] weight gradient of a token GEMM: added in place into the parameters' gradient buffers
    (iit_conv2d_wgrad_tc_scatter; autograd gets None) when `w_direct` holds their table, else returned as strided views."""
    need = int(_lib.lib.iit_conv2d_wgrad_tc_scratch_bytes(C.byref(d)))
    scratch = torch.empty(max(need, 16), dtype=torch.uint8, device=x.device)
    if w_direct is not None:
        table, n_maps, biggest = w_direct
        call("iit_conv2d_wgrad_tc_scatter", _p(g), _p(x), C.byref(d), _p(scratch), need, _p(table), n_maps, biggest, _stream())
        return [None] * len(specs)
    dwb = torch.empty(d.k * d.c, dtype=torch.float32, device=x.device)   # [out_pad][in]
    call("iit_conv2d_wgrad_tc", _p(g), _p(x), _p(dwb), C.byref(d), _p(scratch), need, _stream())
    return [torch.as_strided(dwb, size, stride, offset) for size, stride, offset in specs]


def linear_prepared(x2d, w_oi, w_io, n_in, n_out_pad, specs, params):
    return _LinearPreparedFn.apply(x2d, w_oi, w_io, n_in, n_out_pad, specs, *params)


class GeluLink:
    """Couples the MLP's two linear maps across `mlp.hook_post` so that the GELU's backward runs in a GEMM epilogue:
    the input map (`act=1`) leaves its pre-activation here; the output map's backward then computes
    `dpre = (dy @ W_out^T) * gelu'(pre)` in one kernel (iit_gemm_dgrad_gelu_tc, with the column sums of dpre = the
    input bias's gradient), hands it on as the gradient of `post`, and marks the link; the input map's backward sees the
    mark (same backward pass) and takes the gradient as dpre instead of applying gelu' itself. Anything between the two
    that only zeroes parts of the gradient (a patch hook's backward) commutes with that. `sums` is the site's persistent
    fp64 [2, d_mlp] accumulator (zero between uses)."""
    __slots__ = ("pre", "sums", "state", "_task")

    def __init__(self, sums: torch.Tensor, state: dict):
        """One link per forward; `sums` and `state` ({"dirty": bool}) belong to the site and persist."""
        self.pre, self.sums, self.state, self._task = None, sums, state, -1

    def begin_accumulate(self):
        if self.state.get("dirty"):      # sums left behind by a backward whose gradient never reached the input map
            self.sums.zero_()
        self.state["dirty"] = True

    def mark(self):
        self._task = GradSlot._current_task()

    def marked(self) -> bool:
        t = GradSlot._current_task()
        ok = t != -1 and t == self._task
        self._task = -1
        return ok

    def consumed(self):
        self.state["dirty"] = False


class _LinearEpiFn(torch.autograd.Function):
    """y = act(x @ W + bias (+ resid)) with bias / residual / gelu_new applied to the fp32 accumulator in the GEMM epilogue
    (iit_gemm_bias_act_tc), from the model's PREPARED bf16 operands (see _LinearPreparedFn). `bias`: fp32 [n_out_pad] or
    None, its gradient (the column sums of dy) is returned for each of `bias_params` as the view its spec names.
    `link_out`: this is the MLP's input map (act must be 1); `link_in`: this map consumes that activation (GeluLink)."""

    @staticmethod
    def forward(ctx, x, w_oi, w_io, n_in, n_out_pad, specs, bias, bias_specs, resid, act, link_out, link_in, n_w, *params):
        # `link_in` may be a (GeluLink or None, GradSlot) pair: the slot receives this map's residual gradient (the skip
        # connection's share of the residual stream's gradient) for the LayerNorm backward to add onto
        ctx.res_slot = None
        if isinstance(link_in, tuple):
            link_in, ctx.res_slot = link_in
        # bias_specs may end with the string key "direct": (specs..., ("direct", buffer)) -- the bias gradient (column sums of
        # dy) is then ADDED straight into `buffer`, the optimizer's fp32 gradient view (direct_grad), and autograd gets None
        ctx.bias_direct = None
        if bias_specs is not None and len(bias_specs) and isinstance(bias_specs[-1], tuple) and bias_specs[-1][0] == "direct":
            ctx.bias_direct = bias_specs[-1][1]
            bias_specs = bias_specs[:-1]
        _require_cuda(x, w_oi, w_io, bias, resid)
        x = x.contiguous()
        T = x.shape[0]
        if x.dtype != torch.bfloat16 or n_in % 64 or n_out_pad % 64:
            raise RuntimeError("linear_epi: bf16 activations and widths that are multiples of 64 only")
        y = torch.empty((T, n_out_pad), dtype=x.dtype, device=x.device)
        want_grad = any(ctx.needs_input_grad)
        pre = torch.empty_like(y) if (act == 1 and want_grad) else None
        if resid is not None:
            resid = resid.contiguous()
        call("iit_gemm_bias_act_tc", _p(x), _p(w_oi), _p(bias), _p(resid), _p(y), _p(pre), T, n_in, n_out_pad, int(act), _stream())
        if link_out is not None:
            link_out.pre = pre
        ctx.save_for_backward(x, w_io, pre)
        ctx.geom = (T, n_in, n_out_pad)
        ctx.specs, ctx.w_direct = _split_direct(specs)
        ctx.bias_specs, ctx.act, ctx.has_res = bias_specs, act, resid is not None
        ctx.link_out, ctx.link_in, ctx.n_w = link_out, link_in, n_w
        return y

    @staticmethod
    def backward(ctx, dy):
        x, w_io, pre = ctx.saved_tensors
        T, n_in, n_out = ctx.geom
        dy = dy.contiguous()
        st = _stream()
        d = Conv()
        d.n, d.h, d.w, d.c, d.k, d.r, d.s, d.stride, d.pad, d.p, d.q = T, 1, 1, n_in, n_out, 1, 1, 1, 0, 1, 1
        db = None
        want_db = ctx.bias_specs is not None and any(ctx.needs_input_grad[13 + ctx.n_w:])
        direct = ctx.bias_direct if (want_db and ctx.act != 1) else None
        if direct is not None:      # column sums added into the optimizer's gradient buffer: no zero fill, no AccumulateGrad
            want_db = False
        if ctx.act == 1:
            link = ctx.link_out
            if link is not None and link.marked():
                g = dy                                       # already dpre: the output map's backward applied gelu'
                if want_db:
                    db = torch.empty(n_out, dtype=torch.float32, device=dy.device)
                    scratch = torch.empty(n_out, dtype=torch.float32, device=dy.device)
                    call("iit_bn_param_grads", _p(link.sums), n_out, _p(db), _p(scratch), 0, 1, st)   # fp64 sums -> fp32, cleared
                else:
                    link.sums.zero_()
                link.consumed()
            else:
                g = torch.empty_like(dy)
                call("iit_gelu_bwd", _p(dy), _p(pre), _p(g), _dt(dy), dy.numel(), st)
                if want_db:
                    db = colsum(g)
        else:
            g = dy
            if direct is not None:
                call("iit_colsum", _p(g), _dt(g), _p(direct), g.shape[0], n_out, st)
            elif want_db:
                db = colsum(g)
        dx = None
        if ctx.needs_input_grad[0]:
            dx = torch.empty_like(x)
            lin = ctx.link_in
            if lin is not None and lin.pre is not None and lin.pre.shape == x.shape:
                lin.begin_accumulate()
                call("iit_gemm_dgrad_gelu_tc", _p(g), _p(w_io), _p(lin.pre), _p(dx), _p(lin.sums[0]), _p(lin.sums[1]), T, n_in, n_out, st)
                lin.mark()
            else:
                call("iit_conv2d_dgrad_tc", _p(g), _p(w_io), _p(dx), C.byref(d), st)
        grads = [None] * ctx.n_w
        if any(ctx.needs_input_grad[13:13 + ctx.n_w]):
            grads = _wgrad_into(g, x, d, ctx.specs, ctx.w_direct)
        if ctx.has_res and ctx.res_slot is not None and USE_GRAD_SLOTS:
            ctx.res_slot.put(g)
        bgrads = []
        if ctx.bias_specs is not None:
            bgrads = [torch.as_strided(db, size, stride, offset) if db is not None else None for size, stride, offset in ctx.bias_specs]
        return (dx, None, None, None, None, None, None, None, (g if ctx.has_res else None), None, None, None, None, *grads, *bgrads)


def linear_epi(x2d, w_oi, w_io, n_in, n_out_pad, specs, params, bias=None, bias_specs=None, bias_params=(), resid=None, act=0,
               link_out=None, link_in=None, res_slot=None):
    if res_slot is not None:
        link_in = (link_in, res_slot)
    return _LinearEpiFn.apply(x2d, w_oi, w_io, n_in, n_out_pad, specs, bias, bias_specs, resid, act, link_out, link_in,
                              len(params), *params, *bias_params)


def colsum(x2d: torch.Tensor) -> torch.Tensor:
    out = torch.zeros(x2d.shape[1], dtype=torch.float32, device=x2d.device)
    call("iit_colsum", _p(x2d), _dt(x2d), _p(out), x2d.shape[0], x2d.shape[1], _stream())
    return out


class _BiasAddFn(torch.autograd.Function):
    """out = x + bias (+ resid)"""

    @staticmethod
    def forward(ctx, x, bias, resid):
        _require_cuda(x, bias, resid)
        x = x.contiguous()
        if resid is not None:
            resid = resid.contiguous()
        out = torch.empty_like(x)
        rows, n = x.shape
        call("iit_bias_act_fwd", _p(x), _p(bias), _p(resid), None, _p(out), _dt(x), rows, n, 0, _stream())
        ctx.has_bias, ctx.has_res = bias is not None, resid is not None
        return out

    @staticmethod
    def backward(ctx, dy):
        dy = dy.contiguous()
        db = colsum(dy) if (ctx.has_bias and ctx.needs_input_grad[1]) else None
        return dy, db, (dy if ctx.has_res else None)


def bias_add(x2d, bias, resid=None):
    return _BiasAddFn.apply(x2d, bias, resid)


class _BiasGeluFn(torch.autograd.Function):
    """post = gelu_new(x + bias), returned in `out_shape` (a fresh tensor, not a view, so a patch hook may
    overwrite it in place). The pre-activation is kept for the backward only."""

    @staticmethod
    def forward(ctx, x, bias, out_shape):
        _require_cuda(x, bias)
        x = x.contiguous()
        pre = torch.empty_like(x)
        post = torch.empty(out_shape, dtype=x.dtype, device=x.device)
        rows, n = x.shape
        call("iit_bias_act_fwd", _p(x), _p(bias), None, _p(pre), _p(post), _dt(x), rows, n, 1, _stream())
        ctx.save_for_backward(pre)
        return post

    @staticmethod
    def backward(ctx, dpost):
        (pre,) = ctx.saved_tensors
        dpost = dpost.contiguous()
        dpre = torch.empty_like(pre)
        call("iit_gelu_bwd", _p(dpost), _p(pre), _p(dpre), _dt(pre), pre.numel(), _stream())
        db = colsum(dpre) if ctx.needs_input_grad[1] else None
        return dpre, db, None


def bias_gelu(x2d, bias, out_shape):
    return _BiasGeluFn.apply(x2d, bias, tuple(out_shape))


class _BiasGeluLeanFn(torch.autograd.Function):
    """post = gelu_new(h + bias) for bf16 activations, keeping `h` (the GEMM's raw result, alive anyway) instead of writing
    the pre-activation: the forward moves two tensors instead of three, and the backward computes
    dpre = dpost * gelu'(h + bias) together with the bias gradient (column sums of dpre) in ONE pass
    (iit_gelu_bwd_bias_colsum) instead of a GELU-backward kernel plus a column-sum kernel."""

    @staticmethod
    def forward(ctx, h, bias, out_shape, bias_grad=None):
        _require_cuda(h, bias)
        h = h.contiguous()
        ctx.bias_grad = bias_grad       # the optimizer's gradient view of `bias` (direct_grad): the column sums are added into it
        if h.dtype != torch.bfloat16 or h.shape[1] % 8:
            raise RuntimeError("bias_gelu_lean: bf16 activations with a width that is a multiple of 8 only")
        post = torch.empty(out_shape, dtype=h.dtype, device=h.device)
        rows, n = h.shape
        call("iit_bias_act_fwd", _p(h), _p(bias), None, None, _p(post), _dt(h), rows, n, 1, _stream())
        ctx.save_for_backward(h, bias)
        return post

    @staticmethod
    def backward(ctx, dpost):
        h, bias = ctx.saved_tensors
        dpost = dpost.contiguous()
        rows, n = h.shape
        dpre = torch.empty_like(h)
        if ctx.needs_input_grad[1] and ctx.bias_grad is not None:
            call("iit_gelu_bwd_bias_colsum", _p(dpost), _p(h), _p(bias), _p(dpre), _p(ctx.bias_grad), rows, n, _stream())
            return dpre, None, None, None
        db = torch.zeros(n, dtype=torch.float32, device=h.device) if ctx.needs_input_grad[1] else None
        call("iit_gelu_bwd_bias_colsum", _p(dpost), _p(h), _p(bias), _p(dpre), _p(db), rows, n, _stream())
        return dpre, db, None, None


def bias_gelu_lean(h2d, bias, out_shape, bias_grad=None):
    return _BiasGeluLeanFn.apply(h2d, bias, tuple(out_shape), bias_grad)


class _LNPreFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, eps, slot):
        _require_cuda(x)
        x = x.contiguous()
        rows, d = x.shape
        y = torch.empty_like(x)
        rstd = torch.empty(rows, dtype=torch.float32, device=x.device)
        call("iit_lnpre_fwd", _p(x), _p(y), _p(rstd), _dt(x), rows, d, float(eps), _stream())
        ctx.save_for_backward(y, rstd)
        ctx.slot = slot
        return y

    @staticmethod
    def backward(ctx, dy):
        y, rstd = ctx.saved_tensors
        dy = dy.contiguous()
        first = ctx.slot.take() if (ctx.slot is not None and USE_GRAD_SLOTS) else None
        if first is not None and first.shape == y.shape and first.dtype == y.dtype and first.is_contiguous():
            # the skip connection's gradient of the same residual-stream tensor is already there: add onto it (GradSlot)
            call("iit_lnpre_bwd_accumulate", _p(dy), _p(y), _p(rstd), _p(first), _dt(y), y.shape[0], y.shape[1], _stream())
            return None, None, None
        dx = torch.empty_like(y)
        call("iit_lnpre_bwd", _p(dy), _p(y), _p(rstd), _p(dx), _dt(y), y.shape[0], y.shape[1], _stream())
        return dx, None, None


def lnpre(x2d, eps, slot=None):
    """`slot`: a GradSlot shared with the skip connection that consumes the same tensor (linear_epi(..., res_slot=slot))."""
    return _LNPreFn.apply(x2d, eps, slot)


class _AttnFn(torch.autograd.Function):
    """qkv [B,S,3,H,DH] -> z [B,S,H,DH], causal softmax(q k^T / sqrt(DH)) v."""

    @staticmethod
    def forward(ctx, qkv):
        _require_cuda(qkv)
        qkv = qkv.contiguous()
        b, s, _, h, dh = qkv.shape
        z = torch.empty((b, s, h, dh), dtype=qkv.dtype, device=qkv.device)
        call("iit_attn_fwd", _p(qkv), _p(z), _dt(qkv), b, s, h, dh, _stream())
        ctx.save_for_backward(qkv)
        return z

    @staticmethod
    def backward(ctx, dz):
        (qkv,) = ctx.saved_tensors
        b, s, _, h, dh = qkv.shape
        dz = dz.contiguous()
        dqkv = torch.empty_like(qkv)
        call("iit_attn_bwd", _p(qkv), _p(dz), _p(dqkv), _dt(qkv), b, s, h, dh, _stream())
        return dqkv


def attention(qkv):
    return _AttnFn.apply(qkv)


class _EmbedFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, tokens, W_E, W_pos, dtype, direct=None):
        _require_cuda(tokens, W_E, W_pos)
        ctx.direct = direct     # (W_E.grad, W_pos.grad) views of the optimizer's buffer: the scatter-add goes straight into them
        tokens = tokens.contiguous()
        b, s = tokens.shape
        v, dm = W_E.shape
        out = torch.empty((b, s, dm), dtype=dtype, device=tokens.device)
        call("iit_embed_fwd", _p(tokens), _p(W_E.contiguous()), _p(W_pos.contiguous()), _p(out), _DT[dtype], b * s, s, dm, v,
             _stream())
        ctx.save_for_backward(tokens)
        ctx.shapes = (W_E.shape, W_pos.shape)
        return out

    @staticmethod
    def backward(ctx, dres):
        (tokens,) = ctx.saved_tensors
        dres = dres.contiguous()
        (v, dm), pshape = ctx.shapes
        b, s = tokens.shape
        if ctx.direct is not None:
            call("iit_embed_bwd", _p(tokens), _p(dres), _dt(dres), _p(ctx.direct[0]), _p(ctx.direct[1]), b * s, s, dm, v, _stream())
            return None, None, None, None, None
        dWE = torch.zeros((v, dm), dtype=torch.float32, device=dres.device)
        dWpos = torch.zeros(pshape, dtype=torch.float32, device=dres.device)
        b, s = tokens.shape
        call("iit_embed_bwd", _p(tokens), _p(dres), _dt(dres), _p(dWE), _p(dWpos), b * s, s, dm, v, _stream())
        return None, dWE, dWpos, None, None


def embed(tokens, W_E, W_pos, dtype):
    direct = None
    if torch.is_grad_enabled():
        gE, gP = direct_grad(W_E), direct_grad(W_pos)
        if gE is not None and gP is not None and gE.is_contiguous() and gP.is_contiguous() and W_E.is_contiguous() and W_pos.is_contiguous():
            direct = (gE, gP)
    return _EmbedFn.apply(tokens, W_E, W_pos, dtype, direct)


# =================================================================================================
# either side of the step: batches of pairs, counterfactual inputs
# =================================================================================================
_err_flags: dict = {}


def _err_flag(device) -> torch.Tensor:
    dev = torch.device(device)
    if dev.type == "cuda" and dev.index is None:
        dev = torch.device("cuda", torch.cuda.current_device())
    f = _err_flags.get(dev)
    if f is None:
        f = _err_flags[dev] = torch.zeros(1, dtype=torch.int32, device=dev)
    return f


def raise_on_index_error(device="cuda"):
    """Synchronising check of the out-of-range flag the gather kernels set instead of faulting."""
    f = _err_flag(device)
    if int(f.item()):
        f.zero_()
        raise IndexError("a gather kernel was given an index outside its source tensor")


def gather_rows(src: torch.Tensor, idx: torch.Tensor) -> torch.Tensor:
    """src[idx] along dim 0 for a dense `src` (any trailing shape) and int64 `idx`, as one coalesced gather."""
    _require_cuda(src, idx)
    if not src.is_contiguous():
        raise ValueError("gather_rows: source must be dense (contiguous)")
    if idx.dtype != torch.int64 or idx.dim() != 1:
        raise ValueError("gather_rows: idx must be a 1-D int64 tensor")
    out = torch.empty((idx.shape[0],) + tuple(src.shape[1:]), dtype=src.dtype, device=src.device)
    row_bytes = (src.numel() // max(src.shape[0], 1)) * src.element_size()
    if idx.shape[0] and row_bytes:
        call("iit_gather_rows", _p(out), _p(src), _p(idx.contiguous()), idx.shape[0], row_bytes, src.shape[0],
             _p(_err_flag(src.device)), _stream())
    return out


def gather_box_(dst: torch.Tensor, bank: torch.Tensor, idx: torch.Tensor, r0: int, c0: int) -> torch.Tensor:
    """In place: dst[b, :, r0:r0+h, c0:c0+w] = bank[idx[b]] for bank [items, C, h, w] (dense)."""
    _require_cuda(dst, bank, idx)
    n, c, hh, ww = dst.shape
    items, cb, h, w = bank.shape
    if cb != c or dst.dtype != bank.dtype or not bank.is_contiguous() or r0 + h > hh or c0 + w > ww or idx.shape[0] != n:
        raise ValueError("gather_box_: shape / dtype mismatch")
    strides = (C.c_int64 * 4)(*dst.stride())
    call("iit_gather_box", _p(dst), strides, _p(bank), _p(idx.contiguous()), _dt(dst), n, c, h, w, int(r0), int(c0), items,
         _p(_err_flag(dst.device)), _stream())
    return dst


def pick_other_class(cur_class: torch.Tensor, class_start: torch.Tensor, bank_items: int, seed: int):
    """(bank index, its label) per sample, uniform over the bank items of every class but cur_class[b]."""
    _require_cuda(cur_class, class_start)
    n = cur_class.shape[0]
    out_idx = torch.empty(n, dtype=torch.int64, device=cur_class.device)
    out_lab = torch.empty(n, dtype=torch.int64, device=cur_class.device)
    call("iit_pick_other_class", _p(cur_class.contiguous()), _p(class_start), class_start.numel() - 1, int(bank_items),
         C.c_uint64(seed & 0xFFFFFFFFFFFFFFFF), n, _p(out_idx), _p(out_lab), _p(_err_flag(cur_class.device)), _stream())
    return out_idx, out_lab


# =================================================================================================
# optimiser pieces
# =================================================================================================
def sqnorm_(g_flat: torch.Tensor, out: torch.Tensor):
    call("iit_sqnorm", _p(g_flat), g_flat.numel(), _p(out), _stream())


def adam_step_dev_(p, g, m, v, lr_dev, beta1, beta2, eps, step_dev, sqnorm, max_norm):
    call("iit_adam_step_dev", _p(p), _p(g), _p(m), _p(v), p.numel(), _p(lr_dev), float(beta1), float(beta2), float(eps),
         _p(step_dev), _p(sqnorm), float(max_norm or 0.0), _stream())


def adam_step_(p, g, m, v, lr, beta1, beta2, eps, step, sqnorm, max_norm):
    call("iit_adam_step", _p(p), _p(g), _p(m), _p(v), p.numel(), float(lr), float(beta1), float(beta2), float(eps),
         int(step), _p(sqnorm), float(max_norm or 0.0), _stream())
