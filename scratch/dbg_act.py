import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import restate as R
from iit_b200.model_pairs import IITBehaviorModelPair
from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
HOOK = "mod.layer3.mod.1.mod.conv2.hook_point"
MODE, NODE = sys.argv[1], sys.argv[2]
PTS = ["mod.layer3.mod.1.mod.conv2.hook_point", "mod.layer3.mod.1.hook_point", "mod.layer4.mod.0.mod.conv1.hook_point", "mod.layer4.mod.0.hook_point",
       "mod.layer4.mod.1.mod.conv1.hook_point", "mod.layer4.mod.1.mod.conv2.hook_point", "mod.layer4.hook_point", "mod.fc.hook_point"]
ll, hl, corr = get_alignment("mnist_pvr", {"mode": MODE, "input_shape": (1, 3, 56, 56)})
ll.load_state_dict(R.resnet18_state_dict(0))
pair = IITBehaviorModelPair(hl, ll, corr, training_args={"lr": 1e-3})
ll.train()
base, src = R.synthetic_pvr(8, 100), R.synthetic_pvr(8, 200)
dev = lambda x: tuple(t.cuda() for t in x)
node = [k for k in corr if k.name == NODE][0]
fw, bw = {}, {}
def mk(name, store_f, store_b):
    def h(t, hook=None):
        store_f[name] = t.detach().clone()
        if t.requires_grad:
            t.register_hook(lambda g: store_b.__setitem__(name, g.detach().clone()))
        return None
    return h
# mine: hooks added AFTER the ablation hook at HOOK (so they see the patched tensor)
hl_out = pair.hl_intervention(dev(base), dev(src), node)
pair.ll_source_run(src[0].cuda(), corr[node])
hooks = [(n.name, pair.make_ll_ablation_hook(n)) for n in corr[node]] + [(p, mk(p, fw, bw)) for p in PTS]
out = ll.run_with_hooks(base[0].cuda(), fwd_hooks=hooks)
loss = pair.loss_fn(out, hl_out); loss.backward()

oll = R.ResNet18Oracle(R.resnet18_state_dict(0))
ocorr = R.pvr_corr(MODE, HOOK, 256, 4)
cache = {}
oll(src[0], cache=cache, cache_filter={HOOK})
ofw, obw = {}, {}
ohooks = {}
for p in PTS:
    cap = mk(p, ofw, obw)
    if p == HOOK:
        name, spec = ocorr[NODE][0]
        def both(t, cap=cap, spec=spec):
            t2 = R.patch(t, cache[HOOK], spec); cap(t2); return t2
        ohooks[p] = both
    else:
        ohooks[p] = (lambda t, cap=cap: (cap(t), t)[1])
oout = oll(base[0], hooks=ohooks)
ol = torch.nn.functional.cross_entropy(oout, R.pvr_hl_intervened(base[2], src[2], NODE)); ol.backward()
print("loss", loss.item(), ol.item())
for p in PTS:
    a, b = fw[p].float().cpu(), ofw[p]
    ef = (a - b).abs().max().item() / (b.abs().max().item() + 1e-30)
    if p in bw and p in obw:
        ga, gb = bw[p].float().cpu(), obw[p]
        eb = (ga - gb).abs().max().item() / (gb.abs().max().item() + 1e-30)
    else:
        eb = float("nan")
    print("%-45s fwd %.2e  bwd %.2e" % (p, ef, eb))
for p in ["mod.layer4.hook_point", "mod.layer4.mod.0.hook_point", "mod.layer3.mod.1.hook_point"]:
    a, b = fw[p].float().cpu(), ofw[p]
    print(p, "mask mismatches", ((a > 0) != (b > 0)).sum().item(), "of", a.numel(), "zeros mine", (a == 0).sum().item(), "oracle", (b == 0).sum().item(),
          "min positive mine", a[a > 0].min().item(), "oracle", b[b > 0].min().item())
p = "mod.layer4.mod.1.mod.conv2.hook_point"
ga, gb = bw[p].float().cpu(), obw[p]
d = (ga - gb).abs()
print("grad diff at", p, "max", d.max().item(), "argmax", torch.nonzero(d == d.max())[0].tolist(), "n big", (d > 0.05 * gb.abs().max()).sum().item())
ch = torch.nonzero(d == d.max())[0][1].item()
print("channel", ch, "mine", ga[:, ch].flatten()[:8].tolist(), "\noracle", gb[:, ch].flatten()[:8].tolist())
print("per-channel err count:", (d.amax(dim=(0, 2, 3)) > 0.01 * gb.abs().max()).sum().item(), "of 512")
x = fw[p].float().cpu()[:, ch]; xo = ofw[p][:, ch]
print("x var mine", x.var(unbiased=False).item(), "oracle", xo.var(unbiased=False).item(), "mean", x.mean().item(), xo.mean().item())
