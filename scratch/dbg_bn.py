import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from oracle import restate as R
from iit_b200 import ops
from iit_b200.model_pairs import IITBehaviorModelPair
from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
MODE, NODE = sys.argv[1], sys.argv[2]
ll, hl, corr = get_alignment("mnist_pvr", {"mode": MODE, "input_shape": (1, 3, 56, 56)})
ll.load_state_dict(R.resnet18_state_dict(0))
pair = IITBehaviorModelPair(hl, ll, corr, training_args={"lr": 1e-3})
ll.train()
orig = ops._BNActFn.backward
cnt = [0]
def checked(ctx, dy):
    x, y, gamma, mean, invstd = ctx.saved_tensors
    out = orig(ctx, dy)
    dyf = dy.float()
    g = dyf * (y > 0) if ctx.relu else dyf
    xh = (x.float() - mean.view(1, -1, 1, 1)) * invstd.view(1, -1, 1, 1)
    dg = (g.double() * xh.double()).sum(dim=(0, 2, 3)); db = g.double().sum(dim=(0, 2, 3))
    e1 = (out[1].double() - dg).abs().max().item() / (dg.abs().max().item() + 1e-30)
    e2 = (out[2].double() - db).abs().max().item() / (db.abs().max().item() + 1e-30)
    n = x.shape[0] * x.shape[2] * x.shape[3]
    dxr = gamma.view(1, -1, 1, 1) * invstd.view(1, -1, 1, 1) * (g - (db / n).float().view(1, -1, 1, 1) - xh * (dg / n).float().view(1, -1, 1, 1))
    e3 = (out[0] - dxr).abs().max().item() / (dxr.abs().max().item() + 1e-30)
    print("bn bwd #%d shape %s relu %d res %d  dgamma %.2e dbeta %.2e dx %.2e  dy_cl %s |dy|max %.3e" % (cnt[0], tuple(x.shape), ctx.relu, ctx.has_res, e1, e2, e3, dy.is_contiguous(memory_format=torch.channels_last), dy.abs().max().item()))
    cnt[0] += 1
    return out
ops._BNActFn.backward = staticmethod(checked)
base, src = R.synthetic_pvr(8, 100), R.synthetic_pvr(8, 200)
dev = lambda x: tuple(t.cuda() for t in x)
node = [k for k in corr if k.name == NODE][0]
loss = pair.get_IIT_loss_over_batch(dev(base), dev(src), node, pair.loss_fn)
loss.backward()
