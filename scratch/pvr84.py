"""The strict train step at 84x84 (batch 2048) with cudaProfilerStart/Stop around two replayed steps, for an ncu launch
list (development tool): ncu --profile-from-start off --metrics gpu__time_duration.sum --csv python scratch/pvr84.py"""
import sys, torch
sys.path.insert(0, ".")
from iit_b200.model_pairs import StrictIITModelPair
from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
from oracle import restate as R
dev = torch.device("cuda", 0); B = int(sys.argv[1]) if len(sys.argv) > 1 else 2048
ll, hl, corr = get_alignment("mnist_pvr", {"input_shape": (1, 3, 84, 84)}, device=dev, dtype=torch.bfloat16)
ll.load_state_dict(R.resnet18_state_dict(0))
pair = StrictIITModelPair(hl, ll, corr, training_args={"lr": 1e-3, "cuda_graphs": True})
opt = pair.make_optimizer(); ll.train()
def gen(seed):
    parts = [R.synthetic_pvr(min(512, B - i), seed * 1000 + i, size=84) for i in range(0, B, 512)]
    return tuple(torch.cat([p[j] for p in parts]).to(dev) for j in range(3))
base, src = gen(51), gen(52)
pair.prepare_graphs(base, src, pair.loss_fn, opt)
for _ in range(2): pair.run_train_step(base, src, pair.loss_fn, opt)
torch.cuda.synchronize(); torch.cuda.profiler.start()
s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
s.record()
for _ in range(2): pair.run_train_step(base, src, pair.loss_fn, opt)
e.record(); torch.cuda.synchronize(); torch.cuda.profiler.stop()
print("ms/step", s.elapsed_time(e) / 2)
