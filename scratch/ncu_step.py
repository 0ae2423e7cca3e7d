"""One eager strict train step (5 LL forwards, 3 backwards, 3 updates) at the bench configuration, bracketed by
cudaProfilerStart/Stop for `ncu --profile-from-start off`. `--ioi` profiles the IOI strict step instead.
usage: python scratch/ncu_step.py [--batch 4096] [--ioi]"""
import argparse
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=4096)
ap.add_argument("--ioi", action="store_true")
args = ap.parse_args()

from oracle import restate as R  # noqa: E402

dev = torch.device("cuda", 0)
if args.ioi:
    from iit_b200.model_pairs import IOI_ModelPair
    from iit_b200.tasks.ioi import IOI_HL, corr, make_ll_model
    ll = make_ll_model(device=dev, dtype=torch.bfloat16)
    (bx, by, bv), names = R.synthetic_ioi(args.batch, 21, one_hot=False)
    (sx, sy, sv), _ = R.synthetic_ioi(args.batch, 22, one_hot=False)
    pair = IOI_ModelPair(IOI_HL(50257, names).to(dev), ll, corr, training_args={"lr": 1e-3, "batch_size": args.batch, "lr_scheduler": None})
    base, src = tuple(t.to(dev) for t in (bx, by, bv)), tuple(t.to(dev) for t in (sx, sy, sv))
else:
    from iit_b200.model_pairs import StrictIITModelPair
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    ll, hl, corr = get_alignment("mnist_pvr", {"input_shape": (1, 3, 56, 56)}, device=dev, dtype=torch.bfloat16)
    ll.load_state_dict(R.resnet18_state_dict(0))
    pair = StrictIITModelPair(hl, ll, corr, training_args={"lr": 1e-3})

    def gen(seed):
        parts = [R.synthetic_pvr(min(512, args.batch - i), seed * 1000 + i) for i in range(0, args.batch, 512)]
        return tuple(torch.cat([p[j] for p in parts]).to(dev) for j in range(3))
    base, src = gen(1), gen(2)
opt = pair.make_optimizer()
ll.train()
for _ in range(2):
    pair.run_train_step(base, src, pair.loss_fn, opt)
torch.cuda.synchronize()
torch.cuda.profiler.start()
m = pair.run_train_step(base, src, pair.loss_fn, opt)
torch.cuda.synchronize()
torch.cuda.profiler.stop()
print(m)
