"""Fused stem tail (bn -> relu -> maxpool) entry points timed alone at the bench shape [N,28,28,64] bf16 (development tool)."""
import ctypes as C
import sys

import torch

sys.path.insert(0, ".")
from iit_b200 import _lib  # noqa: E402

N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
dev = "cuda"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
p = lambda t: None if t is None else C.c_void_p(t.data_ptr())
st = lambda: C.c_void_p(torch.cuda.current_stream().cuda_stream)


def timeit(fn, iters=7):
    fn(); torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2] * 1e3


h = w = 28; c = 64; P = Q = 14
x = torch.randn(N, h, w, c, device=dev).bfloat16()
y = torch.empty(N, P, Q, c, device=dev, dtype=torch.bfloat16)
xwin = torch.empty_like(y)
idx = torch.empty(N, P, Q, c, device=dev, dtype=torch.int8)
dy = torch.randn(N, P, Q, c, device=dev).bfloat16()
dx = torch.empty_like(x)
mean = torch.zeros(c, device=dev); invstd = torch.ones(c, device=dev); gamma = torch.ones(c, device=dev); beta = torch.zeros(c, device=dev)
red = torch.zeros(2, c, dtype=torch.float64, device=dev)
full, pooled = x.numel() * 2, y.numel() * 2
t = timeit(lambda: _lib.call("iit_bn_relu_maxpool_fwd", p(x), p(y), p(idx), None, 1, N, h, w, c, p(mean), p(invstd), p(gamma), p(beta), st()))
print(f"fwd (no xwin) {t:7.1f} us  {(full + pooled * 1.5) / t / 1e3:6.0f} GB/s")
t = timeit(lambda: _lib.call("iit_bn_relu_maxpool_fwd", p(x), p(y), p(idx), p(xwin), 1, N, h, w, c, p(mean), p(invstd), p(gamma), p(beta), st()))
print(f"fwd (xwin)    {t:7.1f} us  {(full + pooled * 2.5) / t / 1e3:6.0f} GB/s")
t = timeit(lambda: _lib.call("iit_bn_relu_maxpool_bwd", p(dy), p(idx), p(x), p(y), p(xwin), p(dx), 1, N, h, w, c, p(mean), p(invstd), p(gamma),
                             p(beta), p(red[0]), p(red[1]), st()))
print(f"bwd (sums + dx) {t:7.1f} us  {(pooled * 4.5 + 2 * full) / t / 1e3:6.0f} GB/s")
