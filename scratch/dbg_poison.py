import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import restate as R
from iit_b200 import ops
from iit_b200.model_pairs import IITBehaviorModelPair
from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
ll, hl, corr = get_alignment("mnist_pvr", {"mode": "q", "input_shape": (1, 3, 56, 56)})
ll.load_state_dict(R.resnet18_state_dict(0))
pair = IITBehaviorModelPair(hl, ll, corr, training_args={"lr": 1e-3})
ll.train()
base, src = R.synthetic_pvr(8, 100), R.synthetic_pvr(8, 200)
dev = lambda x: tuple(t.cuda() for t in x)
node = [k for k in corr if k.name == "hook_br"][0]

def poison():
    # fill the allocator's cached blocks with NaN patterns: small and large pools
    bufs = [torch.full((n,), float("nan"), device="cuda") for n in (64, 256, 1024, 4096, 16384, 65536, 262144, 1 << 20, 1 << 22, 1 << 24) for _ in range(8)]
    del bufs

# trace every op output for NaN
orig_call = ops.call
names = []
def traced(name, *a):
    names.append(name)
    orig_call(name, *a)
ops.call = traced

def run():
    ll.zero_grad(set_to_none=True)
    loss = pair.get_IIT_loss_over_batch(dev(base), dev(src), node, pair.loss_fn)
    loss.backward()
    torch.cuda.synchronize()
    return loss.item(), {k: p.grad.detach().clone() for k, p in ll.named_parameters()}

l0, g0 = run()
for it in range(6):
    poison()
    l, g = run()
    bad = [k for k in g if not torch.isfinite(g[k]).all()]
    diffs = {k: ((g[k] - g0[k]).abs().max() / (g0[k].abs().max() + 1e-30)).item() for k in g}
    worst = max(diffs, key=diffs.get)
    print(f"iter {it}: loss {l:.7f} (first {l0:.7f}) nan-params {len(bad)} first-bad {bad[-1] if bad else None} worst-diff {worst} {diffs[worst]:.2e}")
    if bad:
        print("   all bad:", bad)
