import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import restate as R
from iit_b200.model_pairs import IITBehaviorModelPair
from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
HOOK = "mod.layer3.mod.1.mod.conv2.hook_point"
torch.set_num_threads(8)
MODE=sys.argv[1]; NODE=sys.argv[2]
ll, hl, corr = get_alignment("mnist_pvr", {"mode": MODE, "input_shape": (1, 3, 56, 56)})
ll.load_state_dict(R.resnet18_state_dict(0))
pair = IITBehaviorModelPair(hl, ll, corr, training_args={"lr": 1e-3})
ll.train()
base, src = R.synthetic_pvr(8, 100), R.synthetic_pvr(8, 200)
dev = lambda x: tuple(t.cuda() for t in x)
node = [k for k in corr if k.name == NODE][0]
loss = pair.get_IIT_loss_over_batch(dev(base), dev(src), node, pair.loss_fn)
loss.backward()
mine = {k: p.grad.detach().cpu().clone() for k, p in ll.named_parameters()}
# second run, same inputs: determinism check
ll.load_state_dict(R.resnet18_state_dict(0)); ll.zero_grad()
loss2 = pair.get_IIT_loss_over_batch(dev(base), dev(src), node, pair.loss_fn); loss2.backward()
mine2 = {k: p.grad.detach().cpu().clone() for k, p in ll.named_parameters()}
print("run-to-run max rel diff:", max(((mine[k]-mine2[k]).abs().max()/(mine[k].abs().max()+1e-30)).item() for k in mine))
torch.backends.cudnn.allow_tf32 = False; torch.backends.cuda.matmul.allow_tf32 = False

def oracle_grads(device):
    oll = R.ResNet18Oracle(R.resnet18_state_dict(0))
    if device != "cpu":
        oll.p = {k: v.detach().to(device).requires_grad_(True) for k, v in oll.p.items()}
        oll.b = {k: v.to(device) for k, v in oll.b.items()}
    opair = R.OraclePair(oll, "pvr", R.pvr_corr(MODE, HOOK, 256, 4), lr=1e-3)
    b = tuple(t.to(device) for t in base); s = tuple(t.to(device) for t in src)
    # patch restate for device: class map tensor on cpu -> do HL on cpu
    llo, _ = R.ll_intervention(oll, b[0], s[0], opair.corr[NODE])
    hlo = R.pvr_hl_intervened(base[2], src[2], NODE).to(device)
    l = torch.nn.functional.cross_entropy(llo, hlo)
    l.backward()
    return l.item(), {k: v.grad.detach().cpu() for k, v in oll.p.items()}

lc, gc = oracle_grads("cpu")
lg, gg = oracle_grads("cuda")
print("loss mine %.7f cpu %.7f torch-gpu %.7f" % (loss.item(), lc, lg))
print("%-55s %10s %10s %10s" % ("param", "|g|max", "mine-cpu", "tgpu-cpu"))
for k in mine:
    sc = gc[k].abs().max().item()
    e1 = (mine[k] - gc[k]).abs().max().item() / (sc + 1e-30)
    e2 = (gg[k] - gc[k]).abs().max().item() / (sc + 1e-30)
    flag = " <<<" if e1 > 10 * max(e2, 1e-5) else ""
    print("%-55s %10.3e %10.3e %10.3e%s" % (k, sc, e1, e2, flag))
