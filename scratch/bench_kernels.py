"""Per-kernel timings at the bench batch (4096 x 56x56): every ResNet-18 convolution shape (fprop / dgrad / wgrad),
BN forward / backward, pools. CUDA events, L2 flushed between iterations. Development tool, not a test."""
import ctypes as C
import sys

import torch

sys.path.insert(0, ".")
from iit_b200 import _lib, ops  # noqa: E402
from iit_b200._lib import Conv  # noqa: E402

dev = "cuda"
N = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
which = sys.argv[2] if len(sys.argv) > 2 else "all"
flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)


def timeit(fn, iters=5):
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        fn()
        e.record()
        torch.cuda.synchronize()
        ts.append(s.elapsed_time(e))
    ts.sort()
    return ts[len(ts) // 2] * 1e-3


def p(t):
    return C.c_void_p(t.data_ptr())


def st():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


shapes = [  # (name, C, K, H, W, r, stride, pad)
    ("stem1x1", 192, 64, 28, 28, 1, 1, 0),
    ("l1", 64, 64, 14, 14, 3, 1, 1),
    ("l2.0.c1", 64, 128, 14, 14, 3, 2, 1),
    ("l2.ds", 64, 128, 14, 14, 1, 2, 0),
    ("l2", 128, 128, 7, 7, 3, 1, 1),
    ("l3.0.c1", 128, 256, 7, 7, 3, 2, 1),
    ("l3.ds", 128, 256, 7, 7, 1, 2, 0),
    ("l3", 256, 256, 4, 4, 3, 1, 1),
    ("l4.0.c1", 256, 512, 4, 4, 3, 2, 1),
    ("l4.ds", 256, 512, 4, 4, 1, 2, 0),
    ("l4", 512, 512, 2, 2, 3, 1, 1),
]

if which in ("all", "conv"):
    print(f"{'conv':10s} {'GFLOP':>7s} | {'fprop us':>9s} {'TF/s':>6s} | {'dgrad us':>9s} {'TF/s':>6s} | {'wgrad us':>9s} {'TF/s':>6s} | min HBM us")
    for name, c, k, h, w, r, stride, pad in shapes:
        d = Conv()
        d.n, d.h, d.w, d.c, d.k, d.r, d.s, d.stride, d.pad = N, h, w, c, k, r, r, stride, pad
        d.p = (h + 2 * pad - r) // stride + 1
        d.q = (w + 2 * pad - r) // stride + 1
        x = torch.randn(N, h, w, c, device=dev).bfloat16()
        wt = (torch.randn(k, r, r, c, device=dev) * 0.05).bfloat16()
        wt2 = (torch.randn(c, r, r, k, device=dev) * 0.05).bfloat16()
        y = torch.empty(N, d.p, d.q, k, device=dev, dtype=torch.bfloat16)
        dx = torch.empty_like(x)
        dw = torch.empty(k, r, r, c, device=dev)
        sums = torch.zeros(2, k, dtype=torch.float64, device=dev)
        need = int(_lib.lib.iit_conv2d_wgrad_tc_scratch_bytes(C.byref(d)))
        scratch = torch.empty(need, dtype=torch.uint8, device=dev)
        fl = 2.0 * N * d.p * d.q * k * r * r * c
        tf = timeit(lambda: _lib.call("iit_conv2d_fprop_tc", p(x), p(wt), p(y), C.byref(d), p(sums[0]), p(sums[1]), st()))
        tdg = timeit(lambda: _lib.call("iit_conv2d_dgrad_tc", p(y), p(wt2), p(dx), C.byref(d), st()))
        tw = timeit(lambda: _lib.call("iit_conv2d_wgrad_tc", p(y), p(x), p(dw), C.byref(d), p(scratch), need, st()))
        hbm = (x.numel() + y.numel()) * 2 / 6.463e12
        print(f"{name:10s} {fl / 1e9:7.1f} | {tf * 1e6:9.1f} {fl / tf / 1e12:6.0f} | {tdg * 1e6:9.1f} {fl / tdg / 1e12:6.0f} | "
              f"{tw * 1e6:9.1f} {fl / tw / 1e12:6.0f} | {hbm * 1e6:7.1f}")

if which in ("all", "conv", "stem"):
    img = torch.rand(N, 3, 56, 56, device=dev)
    s2d = torch.empty(N, 31, 31, 16, dtype=torch.bfloat16, device=dev)
    w7 = torch.randn(64, 3, 7, 7, device=dev) * 0.1
    w2 = torch.empty(64, 4, 4, 16, dtype=torch.bfloat16, device=dev)
    y = torch.empty(N, 28, 28, 64, dtype=torch.bfloat16, device=dev)
    sums = torch.zeros(2, 64, dtype=torch.float64, device=dev)
    _lib.call("iit_stem_filter_s2d", p(w7), (C.c_int64 * 4)(*w7.stride()), p(w2), st())
    t1 = timeit(lambda: _lib.call("iit_stem_s2d", p(img), _lib.IIT_F32, (C.c_int64 * 4)(*img.stride()), p(s2d), N, 56, 56, st()))
    t2 = timeit(lambda: _lib.call("iit_stem_conv_s2d", p(s2d), p(w2), p(y), N, 31, 31, p(sums[0]), p(sums[1]), st()))
    fl = 2.0 * N * 784 * 64 * 147
    print(f"stem s2d transform {t1 * 1e6:8.1f} us ({(img.numel() * 4 + s2d.numel() * 2) / t1 / 1e9:6.0f} GB/s)   "
          f"stem conv (shifted window) {t2 * 1e6:8.1f} us  {fl / t2 / 1e12:5.0f} TF/s (147-tap FLOPs)  "
          f"{(s2d.numel() * 2 + y.numel() * 2) / t2 / 1e9:6.0f} GB/s")

if which in ("all", "bn"):
    print("\nBN / pool kernels (bf16): us and GB/s of algorithmic bytes")
    for name, c, hw in [("stem", 64, 28), ("l1", 64, 14), ("l2", 128, 7), ("l3", 256, 4), ("l4", 512, 2)]:
        x = torch.randn(N, c, hw, hw, device=dev).bfloat16().contiguous(memory_format=torch.channels_last)
        res = torch.randn_like(x)
        g = torch.ones(c, device=dev)
        b = torch.zeros(c, device=dev)
        rm, rv = torch.zeros(c, device=dev), torch.ones(c, device=dev)
        xr = x.detach().clone().requires_grad_(True)
        rr = res.detach().clone().requires_grad_(True)
        gr = g.clone().requires_grad_(True)
        br = b.clone().requires_grad_(True)
        nbytes = x.numel() * 2

        def fwd():
            return ops.bn_act(xr, gr, br, rr, rm, rv, True, 0.1, 1e-5, True, None)
        yv = fwd()
        dy = torch.randn_like(yv)
        t_f = timeit(fwd)
        t_fb = timeit(lambda: fwd().backward(dy))
        print(f"bn+res+relu {name:5s} fwd(stats+apply) {t_f * 1e6:8.1f} us ({4 * nbytes / t_f / 1e9:6.0f} GB/s of 4 passes)  "
              f"bwd(reduce+apply) {(t_fb - t_f) * 1e6:8.1f} us ({8 * nbytes / (t_fb - t_f) / 1e9:6.0f} GB/s of 8 passes)")
    x = torch.randn(N, 64, 28, 28, device=dev).bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    t_f = timeit(lambda: ops.maxpool3x3s2(x))
    yv = ops.maxpool3x3s2(x)
    dy = torch.randn_like(yv)
    t_fb = timeit(lambda: ops.maxpool3x3s2(x).backward(dy))
    nb = x.numel() * 2
    print(f"maxpool fwd {t_f * 1e6:8.1f} us ({1.25 * nb / t_f / 1e9:6.0f} GB/s)  bwd {(t_fb - t_f) * 1e6:8.1f} us ({1.25 * nb / (t_fb - t_f) / 1e9:6.0f} GB/s)")
    gam = torch.ones(64, device=dev, requires_grad=True)
    bet = torch.zeros(64, device=dev, requires_grad=True)
    rm, rv = torch.zeros(64, device=dev), torch.ones(64, device=dev)
    t_f = timeit(lambda: ops.bn_relu_maxpool(x, gam, bet, rm, rv, True, 0.1, 1e-5))
    t_fb = timeit(lambda: ops.bn_relu_maxpool(x, gam, bet, rm, rv, True, 0.1, 1e-5).backward(dy))
    print(f"fused stats+bn+relu+pool fwd {t_f * 1e6:8.1f} us ({2.375 * nb / t_f / 1e9:6.0f} GB/s)  bwd {(t_fb - t_f) * 1e6:8.1f} us ({3.75 * nb / (t_fb - t_f) / 1e9:6.0f} GB/s)")
    img = torch.rand(N, 3, 56, 56, device=dev)
    t_i = timeit(lambda: ops.stem_im2col(img))
    print(f"stem im2col {t_i * 1e6:8.1f} us ({(img.numel() * 4 + N * 784 * 192 * 2) / t_i / 1e9:6.0f} GB/s)")
    xx = torch.randn(N, 512, device=dev).requires_grad_(True)
    wl = torch.randn(10, 512, device=dev).requires_grad_(True)
    bl = torch.zeros(10, device=dev).requires_grad_(True)
    dyl = torch.randn(N, 10, device=dev)
    t_f = timeit(lambda: ops.linear(xx, wl, bl))
    t_fb = timeit(lambda: ops.linear(xx, wl, bl).backward(dyl))
    print(f"linear fwd {t_f * 1e6:8.1f} us  bwd {(t_fb - t_f) * 1e6:8.1f} us")
