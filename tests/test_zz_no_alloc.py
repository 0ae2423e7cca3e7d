"""The allocation-free weight-gradient entry point (iit_conv2d_wgrad_noalloc).

It exists for callers of the C-ABI that cannot supply scratch memory because they are not allowed to, or
not able to, allocate on the device (an FFI plug-in that only forwards its caller's tensors, a process that
reserved the whole GPU at start-up and must not cudaMalloc afterwards). The test drives it exactly as such
a caller would -- raw pointers through ctypes, no scratch argument -- and checks it against the ordinary
entry point, which takes caller-provided scratch, on two calls in a row: the second call reuses the
built-in buffer with the first call's partial sums still in it.

Collected last on purpose (file name): a run that stops at the first failure has then run everything else."""
import ctypes as C
import math

import pytest
import torch

pytestmark = pytest.mark.gpu


def _desc(n, c, hw, k, ks, stride, pad):
    from iit_b200._lib import Conv
    d = Conv()
    d.n, d.h, d.w, d.c, d.k, d.r, d.s, d.stride, d.pad = n, hw, hw, c, k, ks, ks, stride, pad
    d.p = d.q = (hw + 2 * pad - ks) // stride + 1
    return d


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("cfg", [(64, 64, 7, 3, 1, 1, 64), (64, 128, 14, 3, 2, 1, 33)])
def test_noalloc_wgrad_matches_scratch_entry_point_twice(cfg, dtype):
    """Both input types the entry point takes (fp32 and bf16 activations / output gradients; the weight gradient and the
    partial sums are fp32 either way)."""
    from iit_b200 import _lib
    c, k, hw, ks, stride, pad, n = cfg
    d = _desc(n, c, hw, k, ks, stride, pad)
    need = int(_lib.lib.iit_conv2d_wgrad_scratch_bytes(C.byref(d)))
    assert need > 0, "the case must exercise the split-K scratch"
    code = _lib.IIT_F32 if dtype == torch.float32 else _lib.IIT_BF16
    g = torch.Generator().manual_seed(21)
    for call_no in range(2):  # second call: the built-in buffer still holds the first call's partial sums
        x = torch.randn(n, hw, hw, c, generator=g).cuda().to(dtype)            # NHWC
        dy = (torch.randn(n, d.p, d.q, k, generator=g).cuda() / math.sqrt(n * d.p * d.q)).to(dtype)
        want = torch.empty(k, ks, ks, c, device="cuda")
        scratch = torch.empty(need, dtype=torch.uint8, device="cuda")
        _lib.call("iit_conv2d_wgrad", _ptr(dy), _ptr(x), _ptr(want), C.byref(d), code, _ptr(scratch), need, _stream())
        got = torch.full((k, ks, ks, c), float("nan"), device="cuda")
        _lib.call("iit_conv2d_wgrad_noalloc", _ptr(dy), _ptr(x), _ptr(got), C.byref(d), code, _stream())
        torch.cuda.synchronize()
        scale = want.abs().max().item()
        # same tolerance as the convolution tests of this suite (2e-4 of the output scale; both entry points read the
        # same inputs and accumulate in fp32, so the bf16 case meets it too)
        assert (got - want).abs().max().item() <= 2e-4 * scale, (cfg, dtype, call_no)


def test_noalloc_wgrad_rejects_oversized_problems():
    from iit_b200 import _lib
    d = _desc(96, 64, 14, 64, 3, 1, 1)   # 60 pixel splits x 64 x 576 fp32 partial sums: more than the built-in 8 MiB
    assert int(_lib.lib.iit_conv2d_wgrad_scratch_bytes(C.byref(d))) > (8 << 20)
    dummy = torch.zeros(16, device="cuda")
    with pytest.raises(RuntimeError, match="built-in buffer holds"):
        _lib.call("iit_conv2d_wgrad_noalloc", _ptr(dummy), _ptr(dummy), _ptr(dummy), C.byref(d), _lib.IIT_F32, _stream())
