"""Times single convolution launches through the C-ABI (fprop / dgrad on the tensor-core entry points), cold L2 each time.
Usage: python scratch/time_conv.py  -> one line per case: median microseconds, output GB/s, TFLOP/s."""
import ctypes as C
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from iit_b200 import _lib, ops

dev = "cuda"
flush = torch.empty(64 * 1024 * 1024, device=dev)


def t(fn, n=15):
    for _ in range(3):
        fn()
    ms = []
    for _ in range(n):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ms.append(s.elapsed_time(e))
    ms.sort()
    return ms[len(ms) // 2] * 1e3


def case(n, h, c, k, r, stride, stats, which="fprop"):
    pad = 1 if r == 3 else 0
    p = (h + 2 * pad - r) // stride + 1
    d = _lib.Conv()
    d.n, d.h, d.w, d.c, d.k, d.r, d.s, d.stride, d.pad, d.p, d.q = n, h, h, c, k, r, r, stride, pad, p, p
    x = torch.randn(n, h, h, c, device=dev).bfloat16()
    y = torch.randn(n, p, p, k, device=dev).bfloat16()
    w = torch.randn(k * r * r * c, device=dev).bfloat16()
    s1 = torch.zeros(k, dtype=torch.float64, device=dev) if stats else None
    s2 = torch.zeros(k, dtype=torch.float64, device=dev) if stats else None
    st = ops._stream()
    if which == "fprop":
        fn = lambda: ops.call("iit_conv2d_fprop_tc", ops._p(x), ops._p(w), ops._p(y), C.byref(d), ops._p(s1), ops._p(s2), st)
        out_bytes = y.numel() * 2
    else:
        fn = lambda: ops.call("iit_conv2d_dgrad_tc", ops._p(y), ops._p(w), ops._p(x), C.byref(d), st)
        out_bytes = x.numel() * 2
    before = _lib.kernel_variant_counts()
    us = t(fn)
    after = _lib.kernel_variant_counts()
    kern = [k_ for k_ in after if after[k_] != before.get(k_, 0)]
    flops = 2.0 * n * p * p * k * r * r * c
    print(f"{which:6s} n={n} {h}x{h} {c}->{k} {r}x{r}/{stride} stats={int(stats)}: {us:7.1f} us  out {out_bytes / us / 1e3:7.0f} GB/s  "
          f"{flops / us / 1e6:7.1f} TF  {kern}", flush=True)


def wcase(n, h, c, k, r, stride):
    pad = 1 if r == 3 else 0
    p = (h + 2 * pad - r) // stride + 1
    d = _lib.Conv()
    d.n, d.h, d.w, d.c, d.k, d.r, d.s, d.stride, d.pad, d.p, d.q = n, h, h, c, k, r, r, stride, pad, p, p
    x = torch.randn(n, h, h, c, device=dev).bfloat16()
    dy = torch.randn(n, p, p, k, device=dev).bfloat16()
    dw = torch.zeros(k * r * r * c, device=dev)
    need = int(_lib.lib.iit_conv2d_wgrad_tc_scratch_bytes(C.byref(d)))
    scratch = torch.empty(need, dtype=torch.uint8, device=dev)
    st = ops._stream()
    fn = lambda: ops.call("iit_conv2d_wgrad_tc_accumulate", ops._p(dy), ops._p(x), ops._p(dw), C.byref(d), ops._p(scratch), need, st)
    before = _lib.kernel_variant_counts()
    us = t(fn)
    after = _lib.kernel_variant_counts()
    kern = [k_ for k_ in after if after[k_] != before.get(k_, 0)]
    flops = 2.0 * n * p * p * k * r * r * c
    print(f"wgrad  n={n} {h}x{h} {c}->{k} {r}x{r}/{stride}: {us:7.1f} us  in {(x.numel() + dy.numel()) * 2 / us / 1e3:7.0f} GB/s  "
          f"{flops / us / 1e6:7.1f} TF  scratch {need / 1e6:.1f} MB {kern}", flush=True)


if __name__ == "__main__":
    N = 4096
    if len(sys.argv) > 1 and sys.argv[1] == "win":        # layer1's shifted-window kernel
        for stats in (True, False):
            case(N, 14, 64, 64, 3, 1, stats)
        case(N, 14, 64, 64, 3, 1, False, "dgrad")
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "stem":       # the stem convolution on the space-to-depth window kernel
        s2d = torch.randn(N, 31, 31, 16, device=dev).bfloat16()
        w2 = torch.randn(64 * 4 * 4 * 16, device=dev).bfloat16()
        y = torch.empty(N, 28, 28, 64, device=dev).bfloat16()
        st = ops._stream()
        for stats in (True, False):
            s1 = torch.zeros(64, dtype=torch.float64, device=dev) if stats else None
            s2 = torch.zeros(64, dtype=torch.float64, device=dev) if stats else None
            us = t(lambda: ops.call("iit_stem_conv_s2d", ops._p(s2d), ops._p(w2), ops._p(y), N, 31, 31, ops._p(s1), ops._p(s2), st))
            print(f"stem s2d conv n={N} stats={int(stats)}: {us:.1f} us  out {y.numel() * 2 / us / 1e3:.0f} GB/s", flush=True)
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "wintma":     # shifted-window kernels with / without the TMA-store epilogue
        s2d = torch.randn(N, 31, 31, 16, device=dev).bfloat16()
        w2 = torch.randn(64 * 4 * 4 * 16, device=dev).bfloat16()
        y = torch.empty(N, 28, 28, 64, device=dev).bfloat16()
        st = ops._stream()
        s1 = torch.zeros(64, dtype=torch.float64, device=dev); s2 = torch.zeros(64, dtype=torch.float64, device=dev)
        for on in (1, 0, 1):
            _lib.lib.iit_conv_win_tma_store(on)
            print("TMA-store epilogue", on)
            us = t(lambda: ops.call("iit_stem_conv_s2d", ops._p(s2d), ops._p(w2), ops._p(y), N, 31, 31, ops._p(s1), ops._p(s2), st))
            print(f"stem s2d conv n={N} stats=1: {us:.1f} us  out {y.numel() * 2 / us / 1e3:.0f} GB/s", flush=True)
            case(N, 14, 64, 64, 3, 1, True)
            case(N, 14, 64, 64, 3, 1, False, "dgrad")
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "pair128":    # layer2 on CTA pairs (mode 4) against the two-box single-CTA tiles
        for mode in (2, 1, 2, 1):
            _lib.lib.iit_conv_pair_enable(mode)
            print("pair mode", mode)
            case(N, 7, 128, 128, 3, 1, True)
            case(N, 7, 128, 128, 3, 1, False, "dgrad")
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "pair":       # short reductions with and without the CTA-pair kernel
        for on in (1, 0):
            _lib.lib.iit_conv_pair_enable(on)
            print("pair kernel", on)
            case(N, 7, 128, 256, 1, 2, True)
            case(N, 7, 128, 256, 3, 2, True)
            case(N, 4, 256, 512, 3, 2, True)
            case(N, 4, 256, 256, 3, 1, True)
            case(N, 7, 128, 256, 1, 2, False, "dgrad")
        sys.exit(0)
    if len(sys.argv) > 1 and sys.argv[1] == "short":      # the two launches an ncu capture wants
        case(N, 14, 64, 128, 1, 2, True)
        case(N, 14, 64, 128, 1, 2, False)
        case(N, 7, 128, 128, 3, 1, True)
        sys.exit(0)
    for stats in (True, False):
        case(N, 14, 64, 128, 1, 2, stats)
        case(N, 14, 64, 128, 3, 2, stats)
        case(N, 7, 64, 128, 1, 1, stats)        # same output size as the two above, unit stride
        case(N, 7, 128, 256, 1, 2, stats)
        case(N, 7, 128, 256, 3, 2, stats)
        case(N, 4, 256, 512, 1, 2, stats)
        case(N, 4, 256, 512, 3, 2, stats)
    case(N, 14, 64, 128, 1, 2, False, "dgrad")
    case(N, 7, 128, 256, 1, 2, False, "dgrad")
    case(N, 4, 256, 512, 1, 2, False, "dgrad")
    # the transformer's token GEMMs (IOI, 1024 x 16 tokens)
    T = 16384
    for (c, k) in [(64, 192), (64, 64), (64, 3072), (3072, 64)]:
        case(T, 1, c, k, 1, 1, False)
        case(T, 1, c, k, 1, 1, False, "dgrad")
    case(1024, 1, 64, 50304, 1, 1, False)
    case(1024, 1, 64, 50304, 1, 1, False, "dgrad")
