"""Stress: the statistics a convolution's epilogue returns against the sums of the tensor it wrote, many repetitions."""
import ctypes as C, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from iit_b200 import _lib, ops
dev = "cuda"
torch.manual_seed(0)
worst = 0.0
for (n, h, c, k, r, stride) in [(4096, 7, 128, 128, 3, 1), (4096, 14, 64, 128, 1, 2), (4096, 4, 256, 256, 3, 1), (4096, 2, 512, 512, 3, 1),
                                (4096, 4, 256, 512, 1, 2), (512, 7, 128, 256, 3, 2), (4096, 14, 64, 64, 3, 1), (333, 7, 128, 128, 3, 1)]:
    pad = 1 if r == 3 else 0
    p = (h + 2 * pad - r) // stride + 1
    d = _lib.Conv()
    d.n, d.h, d.w, d.c, d.k, d.r, d.s, d.stride, d.pad, d.p, d.q = n, h, h, c, k, r, r, stride, pad, p, p
    x = torch.randn(n, h, h, c, device=dev).bfloat16()
    w = (torch.randn(k * r * r * c, device=dev) * 0.05).bfloat16()
    y = torch.empty(n, p, p, k, device=dev).bfloat16()
    st = ops._stream()
    bad = 0
    for it in range(60):
        s1 = torch.zeros(k, dtype=torch.float64, device=dev); s2 = torch.zeros(k, dtype=torch.float64, device=dev)
        y.fill_(float("nan"))
        ops.call("iit_conv2d_fprop_tc", ops._p(x), ops._p(w), ops._p(y), C.byref(d), ops._p(s1), ops._p(s2), st)
        yf = y.double().reshape(-1, k)
        e1 = float(((s1 - yf.sum(0)).abs() / (yf.abs().sum(0) + 1e-9)).max())
        e2 = float(((s2 - (yf * yf).sum(0)).abs() / ((yf * yf).sum(0) + 1e-9)).max())
        worst = max(worst, e1, e2)
        if max(e1, e2) > 1e-5 or not torch.isfinite(yf).all():
            bad += 1
    print(f"n={n} {h}x{h} {c}->{k} {r}x{r}/{stride}: bad {bad}/60, worst rel so far {worst:.2e}", flush=True)
