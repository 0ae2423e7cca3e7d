"""One launch of each l2/l3/l4 convolution (fprop, dgrad, wgrad) for an ncu capture."""
import ctypes as C, sys, torch
sys.path.insert(0, ".")
from iit_b200 import _lib
from iit_b200._lib import Conv
dev = "cuda"; N = 4096
p = lambda t: C.c_void_p(t.data_ptr())
st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)
for name, c, k, h, w, r, stride, pad in [("l2", 128, 128, 7, 7, 3, 1, 1), ("l3", 256, 256, 4, 4, 3, 1, 1), ("l4", 512, 512, 2, 2, 3, 1, 1),
                                         ("l3.0.c1", 128, 256, 7, 7, 3, 2, 1)]:
    d = Conv(); d.n, d.h, d.w, d.c, d.k, d.r, d.s, d.stride, d.pad = N, h, w, c, k, r, r, stride, pad
    d.p = (h + 2 * pad - r) // stride + 1; d.q = (w + 2 * pad - r) // stride + 1
    x = torch.randn(N, h, w, c, device=dev).bfloat16()
    wt = (torch.randn(k, r, r, c, device=dev) * 0.05).bfloat16(); wt2 = (torch.randn(c, r, r, k, device=dev) * 0.05).bfloat16()
    y = torch.empty(N, d.p, d.q, k, device=dev, dtype=torch.bfloat16); dx = torch.empty_like(x); dw = torch.empty(k, r, r, c, device=dev)
    sums = torch.zeros(2, k, dtype=torch.float64, device=dev)
    need = int(_lib.lib.iit_conv2d_wgrad_tc_scratch_bytes(C.byref(d))); scratch = torch.empty(need, dtype=torch.uint8, device=dev)
    for _ in range(int(sys.argv[1]) if len(sys.argv) > 1 else 2):
        _lib.call("iit_conv2d_fprop_tc", p(x), p(wt), p(y), C.byref(d), p(sums[0]), p(sums[1]), st())
        _lib.call("iit_conv2d_dgrad_tc", p(y), p(wt2), p(dx), C.byref(d), st())
        _lib.call("iit_conv2d_wgrad_tc", p(y), p(x), p(dw), C.byref(d), p(scratch), need, st())
    torch.cuda.synchronize()
print("ok")
