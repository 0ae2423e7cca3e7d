import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import restate as R
from iit_b200.model_pairs import IITBehaviorModelPair
from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
ll, hl, corr = get_alignment("mnist_pvr", {"mode": "q", "input_shape": (1, 3, 56, 56)})
ll.load_state_dict(R.resnet18_state_dict(0))
pair = IITBehaviorModelPair(hl, ll, corr, training_args={"lr": 1e-3})
ll.train()
base, src = R.synthetic_pvr(8, 100), R.synthetic_pvr(8, 200)
dev = lambda x: tuple(t.cuda() for t in x)
node = [k for k in corr if k.name == "hook_br"][0]
loss = pair.get_IIT_loss_over_batch(dev(base), dev(src), node, pair.loss_fn)
loss.backward()
torch.cuda.synchronize()
print("done", loss.item())
