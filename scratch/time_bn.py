"""Times the BatchNorm passes through the C-ABI at the ResNet's four activation sizes, cold L2: microseconds and GB/s of
algorithmic bytes. python scratch/time_bn.py"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from iit_b200 import ops

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


p, call, st = ops._p, ops.call, ops._stream()
# an empty launch as the floor of this timing method
z = torch.zeros(4, device=dev)
print(f"floor (tiny torch kernel between two events): {t(lambda: z.add_(1)):.1f} us")
for rows, c in [(4096 * 196, 64), (4096 * 49, 128), (4096 * 16, 256), (4096 * 4, 512), (4096 * 49, 64)]:
    x = torch.randn(rows, c, device=dev).bfloat16()
    res = torch.randn(rows, c, device=dev).bfloat16()
    y = torch.empty_like(x); dx = torch.empty_like(x); dres = torch.empty_like(x)
    mask = torch.empty(rows * c // 8, dtype=torch.uint8, device=dev)
    mean = torch.zeros(c, device=dev); invstd = torch.ones(c, device=dev); g = torch.ones(c, device=dev); b = torch.zeros(c, device=dev)
    dg = torch.zeros(c, dtype=torch.float64, device=dev); db = torch.zeros(c, dtype=torch.float64, device=dev)
    nb = x.numel() * 2
    cases = {
        "fwd relu+mask      (r1 w1)": (lambda: call("iit_bn_act_fwd", p(x), None, p(y), 1, rows, c, p(mean), p(invstd), p(g), p(b), 1, p(mask), st), 2 * nb),
        "fwd res relu+mask  (r2 w1)": (lambda: call("iit_bn_act_fwd", p(x), p(res), p(y), 1, rows, c, p(mean), p(invstd), p(g), p(b), 1, p(mask), st), 3 * nb),
        "bwd reduce mask    (r2)   ": (lambda: call("iit_bn_act_bwd_reduce", p(y), p(y), p(x), 1, rows, c, p(mean), p(invstd), 1, p(dg), p(db), p(mask), st), 2 * nb),
        "bwd apply mask     (r2 w1)": (lambda: call("iit_bn_act_bwd_apply", p(y), p(y), p(x), p(dx), None, 1, rows, c, p(mean), p(invstd), p(g), p(dg), p(db), 1, p(mask), st), 3 * nb),
        "bwd apply+dres     (r2 w2)": (lambda: call("iit_bn_act_bwd_apply", p(y), p(y), p(x), p(dx), p(dres), 1, rows, c, p(mean), p(invstd), p(g), p(dg), p(db), 1, p(mask), st), 4 * nb),
        "stats              (r1)   ": (lambda: call("iit_bn_stats", p(x), 1, rows, c, p(dg), p(db), st), nb),
    }
    for name, (fn, bytes_) in cases.items():
        us = t(fn)
        print(f"rows={rows:7d} c={c:3d} ({nb / 1e6:6.1f} MB) {name}: {us:6.1f} us  {bytes_ / us / 1e3:6.0f} GB/s", flush=True)
