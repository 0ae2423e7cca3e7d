import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import restate as R
from iit_b200.model_pairs import IITBehaviorModelPair
from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
HOOK = "mod.layer3.mod.1.mod.conv2.hook_point"
torch.set_num_threads(8)
mode = sys.argv[1] if len(sys.argv) > 1 else "c"
ll, hl, corr = get_alignment("mnist_pvr", {"mode": mode, "input_shape": (1, 3, 56, 56)})
ll.load_state_dict(R.resnet18_state_dict(0))
pair = IITBehaviorModelPair(hl, ll, corr, training_args={"lr": 1e-3})
opt = pair.make_optimizer()
ll.train()
base, src = R.synthetic_pvr(8, 100), R.synthetic_pvr(8, 200)
dev = lambda x: tuple(t.cuda() for t in x)
node = pair.sample_hl_name()
loss = pair.get_IIT_loss_over_batch(dev(base), dev(src), node, pair.loss_fn)
opt.zero_grad(); loss.backward()
gm = {k: p.grad.detach().cpu().clone() for k, p in ll.named_parameters()}
sq = opt.grad_sqnorm().item()
opt.step(max_norm=1.0)
pm = {k: p.detach().cpu().clone() for k, p in ll.named_parameters()}
beh = pair.get_behaviour_loss_over_batch(dev(base), pair.loss_fn).item()

oll = R.ResNet18Oracle(R.resnet18_state_dict(0))
opair = R.OraclePair(oll, "pvr", R.pvr_corr(mode, HOOK, 256, 4), lr=1e-3)
n2 = opair.sample_hl_name()
l2 = opair.iit_loss(base, src, n2)
opair.opt.zero_grad(); l2.backward()
go = {k: v.grad.detach().clone() for k, v in oll.p.items()}
tn = torch.nn.utils.clip_grad_norm_(oll.parameters(), 1.0).item()
opair.opt.step()
po = {k: v.detach().clone() for k, v in oll.p.items()}
o = oll(base[0]); beh2 = torch.nn.functional.cross_entropy(o, base[1]).item()
print("node", node, n2, "loss", loss.item(), l2.item(), "norm", sq ** 0.5, tn, "beh", beh, beh2, "rel", abs(beh-beh2)/beh2)
tot_bad = 0
for k in pm:
    d = (pm[k] - po[k]).abs()
    bad = (d > 1e-4).sum().item()
    tot_bad += bad
    if bad:
        idx = d.flatten().argmax()
        print("%-52s n=%8d bad=%7d maxd=%.2e  g_mine=%.3e g_cpu=%.3e" % (k, d.numel(), bad, d.max().item(), gm[k].flatten()[idx].item(), go[k].flatten()[idx].item()))
print("total elements with |dp|>1e-4:", tot_bad)
allg = torch.cat([g.flatten() for g in go.values()]).abs()
for th in (1e-6, 1e-7, 1e-8, 1e-9, 0):
    print("frac |g_clipped| <= %g: %.4f" % (th, (allg <= th).float().mean().item()))
