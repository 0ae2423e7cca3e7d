"""Where does the bf16 path's gradient differ from the fp32 path's? Per-parameter relative L2 error of the IIT-term
gradient at a given batch, bf16 tensor-core path against this library's exact-fp32 path (same weights, same pairs)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import restate as R
from iit_b200.model_pairs import StrictIITModelPair
from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment

B = int(sys.argv[1]) if len(sys.argv) > 1 else 4096
HOOK = "mod.layer3.mod.1.mod.conv2.hook_point"


def gen(seed):
    parts = [R.synthetic_pvr(min(512, B - i), seed * 1000 + i) for i in range(0, B, 512)]
    return tuple(torch.cat([p[j] for p in parts]).cuda() for j in range(3))


b, s = gen(11), gen(12)
grads = {}
for dtype in (torch.float32, torch.bfloat16):
    ll, hl, corr = get_alignment("mnist_pvr", {"mode": "q", "hook_point": HOOK, "input_shape": (1, 3, 56, 56)}, dtype=dtype)
    ll.load_state_dict(R.resnet18_state_dict(0))
    pair = StrictIITModelPair(hl, ll, corr, training_args={"lr": 1e-3})
    opt = pair.make_optimizer()
    ll.train()
    for term in ("iit", "behaviour"):
        opt.zero_grad()
        node = list(corr.keys())[1]
        loss = pair.get_IIT_loss_over_batch(b, s, node, pair.loss_fn) if term == "iit" else pair.get_behaviour_loss_over_batch(b, pair.loss_fn)
        loss.backward()
        grads[(dtype, term)] = ({k: p.grad.detach().double().clone() for k, p in ll.named_parameters()}, loss.item())
for term in ("iit", "behaviour"):
    g32, l32 = grads[(torch.float32, term)]
    g16, l16 = grads[(torch.bfloat16, term)]
    print(f"== {term}: loss fp32 {l32:.6f} bf16 {l16:.6f}")
    num = den = 0.0
    for k in g32:
        e = float((g16[k] - g32[k]).norm()); n = float(g32[k].norm())
        num += e * e; den += n * n
        cos = float((g16[k] * g32[k]).sum() / (g16[k].norm() * g32[k].norm() + 1e-300))
        print(f"{k:55s} |g| {n:.3e} rel {e / max(n, 1e-300):.3e} cos {cos:.5f}")
    print(f"total rel-L2 {(num / den) ** 0.5:.3e}")
