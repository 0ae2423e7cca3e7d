import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from iit_b200 import ops
torch.manual_seed(0)
V, Vp, B = 50257, 50304, 1024
lg = (torch.randn(B, Vp, device="cuda") * 2).bfloat16()
tgt = torch.randint(0, V, (B,), device="cuda")
flush = torch.empty(64 * 1024 * 1024, device="cuda")
def t(fn, n=20):
    for _ in range(3): fn()
    ms = []
    for _ in range(n):
        flush.zero_()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); fn(); e.record(); torch.cuda.synchronize()
        ms.append(s.elapsed_time(e))
    ms.sort()
    return ms[len(ms) // 2] * 1e3
print("padded bf16 fwd us", t(lambda: ops.cross_entropy(lg, tgt, classes=V)))
lg2 = lg[:, :V].contiguous()
print("unpadded bf16 fwd us", t(lambda: ops.cross_entropy(lg2, tgt)))
lg3 = lg.float()
print("fp32 fwd us", t(lambda: ops.cross_entropy(lg3, tgt, classes=V)))
l = lg.clone().requires_grad_(True)
loss = ops.cross_entropy(l, tgt, classes=V)
print("bwd us", t(lambda: torch.autograd.grad(loss, l, retain_graph=True)))
print("argmax us", t(lambda: ops.argmax_rows(lg[:, :V])))
from oracle import restate as R
(bx, _, _), names = R.synthetic_ioi(B, 3, one_hot=False)
(sx, _, _), _ = R.synthetic_ioi(B, 4, one_hot=False)
bx, sx, nd = bx.cuda(), sx.cuda(), names.cuda()
for node in ("hook_duplicate", "hook_name_mover"):
    print(node, "hl standalone us", t(lambda: ops.hl_ioi_last_argmax(bx, sx, node, nd, V)))
    print(node, "hl + ce fused us", t(lambda: ops.hl_cross_entropy(lg, ("ioi", bx, sx, node, nd, V), classes=V)))
