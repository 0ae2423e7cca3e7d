import sys, time, torch
sys.path.insert(0, ".")
from iit_b200 import _lib, ops
from iit_b200.model_pairs import IOI_ModelPair
from iit_b200.tasks.ioi import IOI_HL, corr, make_ll_model
from oracle import restate as R
dev = "cuda"; B = int(sys.argv[1]) if len(sys.argv) > 1 else 1024
for dtype, graphs in ((torch.bfloat16, False), (torch.bfloat16, True)):
    torch.manual_seed(0)
    ll = make_ll_model(device=dev, dtype=dtype)
    (bx, by, bv), names = R.synthetic_ioi(B, 21, one_hot=False)
    (sx, sy, sv), _ = R.synthetic_ioi(B, 22, one_hot=False)
    hl = IOI_HL(50257, names).to(dev)
    pair = IOI_ModelPair(hl, ll, corr, training_args={"lr": 1e-3, "batch_size": B, "lr_scheduler": None, "cuda_graphs": graphs})
    opt = pair.make_optimizer(); ll.train()
    base = tuple(t.to(dev) for t in (bx, by, bv)); src = tuple(t.to(dev) for t in (sx, sy, sv))
    if graphs: pair.prepare_graphs(base, src, pair.loss_fn, opt)
    for _ in range(2): m = pair.run_train_step(base, src, pair.loss_fn, opt)
    torch.cuda.synchronize(); t0 = time.perf_counter()
    for _ in range(3): m = pair.run_train_step(base, src, pair.loss_fn, opt)
    torch.cuda.synchronize(); dt = (time.perf_counter() - t0) / 3
    print(dtype, "graphs", graphs, f"{dt*1e3:.1f} ms/step  {B/dt:.0f} pairs/s", m)
    if graphs: continue
    # per-entry-point time
    rec = {}
    inner = ops.call
    def timed(name, *a):
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record(); inner(name, *a); e.record(); rec.setdefault(name, []).append((s, e))
    ops.call = timed
    pair.run_train_step(base, src, pair.loss_fn, opt); torch.cuda.synchronize()
    ops.call = inner
    tot = {k: sum(s.elapsed_time(e) for s, e in v) for k, v in rec.items()}
    print("  ", sorted(((round(v, 2), k, len(rec[k])) for k, v in tot.items()), reverse=True)[:10], "sum", round(sum(tot.values()), 1))
