"""The small-reduction convolutions (1x1 stride-2 downsample, 3x3 stride-2 block entry) for an ncu capture (development tool)."""
import ctypes as C, sys, torch
sys.path.insert(0, ".")
from iit_b200 import _lib
from iit_b200._lib import Conv
dev = "cuda"; N = 4096
p = lambda t: C.c_void_p(t.data_ptr())
st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
for name, c, k, h, w, r, stride, pad in [("l2.ds", 64, 128, 14, 14, 1, 2, 0), ("l2.0.c1", 64, 128, 14, 14, 3, 2, 1)]:
    d = Conv(); d.n, d.h, d.w, d.c, d.k, d.r, d.s, d.stride, d.pad = N, h, w, c, k, r, r, stride, pad
    d.p = (h + 2 * pad - r) // stride + 1; d.q = (w + 2 * pad - r) // stride + 1
    x = torch.randn(N, h, w, c, device=dev).bfloat16()
    wt = (torch.randn(k, r, r, c, device=dev) * 0.05).bfloat16(); wt2 = (torch.randn(c, r, r, k, device=dev) * 0.05).bfloat16()
    y = torch.empty(N, d.p, d.q, k, device=dev, dtype=torch.bfloat16); dx = torch.empty_like(x)
    sums = torch.zeros(2, k, dtype=torch.float64, device=dev)
    for _ in range(2):
        _lib.call("iit_conv2d_fprop_tc", p(x), p(wt), p(y), C.byref(d), p(sums[0]), p(sums[1]), st())
        _lib.call("iit_conv2d_dgrad_tc", p(y), p(wt2), p(dx), C.byref(d), st())
    torch.cuda.synchronize()
print("ok")
