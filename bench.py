"""bench.py -- IIT train-step throughput (pairs/s) on the configuration BASELINE.json quotes:
MNIST-PVR, ResNet-18 LL, IIT + behaviour + strict-IIT losses, bf16 activations, batch 4096 pairs
per GPU (configs[1]); one "step" = one StrictIITModelPair.run_train_step over one batch of
synthetic (base, source) pairs = 5 LL forwards, 3 backwards, 3 clip+Adam updates.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch B]
                    [--dtype bf16|f32]

N>1 is launched by torchrun (one rank per GPU, NCCL gradient all-reduce, weak scaling).
Prints ONE JSON line on rank 0. `--impl reference` times the oracle's CPU restatement of the
reference step (oracle/restate.py; the Python reference itself needs /root/reference and
transformer_lens, neither of which exists on the GPU box) on the host cores.
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402
import torch  # noqa: E402

HOOK = "mod.layer3.mod.1.mod.conv2.hook_point"
F56 = 258_162_688            # ResNet-18 forward FLOPs per 56x56 image
METRIC = "iit_train_pairs_per_sec"
UNIT = "pairs/s"


def peaks():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        p = json.load(open(path))
        return {"hbm_gbs": p["hbm_gbs"], "tflops": p.get("bf16_tflops_sustained", p["bf16_tflops"]), "src": "measured"}
    return {"hbm_gbs": 6650.0, "tflops": 1400.0, "src": "fallback"}


def ncu_traffic(entry_point):
    """Average DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) of the kernels behind `entry_point`,
    from the committed `ncu --set full` capture of this same command (profiles/ncu_traffic.json); None if not captured."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(path):
        return None
    ent = json.load(open(path)).get(entry_point)
    return None if ent is None else ent.get("dram_bytes_per_launch")


def workload_config(args, world):
    return {
        "workload": f"MNIST-PVR IIT+behaviour+strict train step, ResNet-18 LL, synthetic 56x56 4-digit grids, "
                    f"batch {args.batch} pairs/GPU ({'bf16' if args.dtype == 'bf16' else 'f32'} activations, "
                    f"fp32 master weights/accumulation)",
        "baseline_config": "configs[1]",
        "batch_pairs_per_gpu": args.batch,
        "global_batch_pairs": args.batch * world,
        "image": "3x56x56",
        "hook_point": HOOK,
        "mode": "q",
        "losses": ["iit", "strict", "behaviour"],
        "optimizer": "Adam lr 1e-3, clip_grad_norm 1.0, one update per loss term",
        "parallelism": f"dp{world}",
        "l2": "no flush needed: per-step activation working set is several GB (>> 126 MB L2)",
    }


# --------------------------------------------------------------------------------------------------
# reference arm / cpu baseline: the oracle's CPU restatement of the same step
# --------------------------------------------------------------------------------------------------
def cpu_step_factory(batch):
    from oracle import restate as R
    torch.set_num_threads(os.cpu_count() or 1)
    oll = R.ResNet18Oracle(R.resnet18_state_dict(0))
    strict = [(n, None) for n in R.ResNet18Oracle.conv_hook_points() if n != HOOK]
    pair = R.OraclePair(oll, "pvr", R.pvr_corr("q", HOOK, 256, 4), lr=1e-3, weights=(1.0, 1.0, 1.0), strict_nodes=strict)
    base, src = R.synthetic_pvr(batch, 11), R.synthetic_pvr(batch, 12)

    def step():
        return pair.train_step(base, src, strict=True)

    return step


def cpu_baseline(budget_s=20.0):
    """Bounded sample of the same workload on the host cores (oracle port)."""
    batch = 32
    step = cpu_step_factory(batch)
    step()  # warm-up
    t0 = time.perf_counter()
    n = 0
    while True:
        step()
        n += 1
        if time.perf_counter() - t0 > budget_s or n >= 50:
            break
    dt = time.perf_counter() - t0
    return {"value": batch * n / dt, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
            "sample": f"{n} steps of {batch} pairs (same step: IIT+strict+behaviour, fp32, torch-CPU eager restatement "
                      f"oracle/restate.py), {dt:.1f} s"}


def run_reference(args, rank, world):
    if rank != 0:
        return
    total = max(args.steps + args.warmup, 1)
    # size the per-step sample so the whole run stays within ~2 minutes
    probe = cpu_step_factory(16)
    probe()
    t0 = time.perf_counter()
    probe()
    per_pair = (time.perf_counter() - t0) / 16
    batch = int(max(8, min(256, 120.0 / (total * per_pair))))
    step = cpu_step_factory(batch)
    for _ in range(args.warmup):
        step()
    t0 = time.perf_counter()
    for _ in range(args.steps):
        step()
    dt = time.perf_counter() - t0
    value = batch * args.steps / dt
    cfg = workload_config(args, world)
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / max(args.steps, 1), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port",
                         "sample": f"{args.steps} steps of {batch} pairs each on the host cores (oracle/restate.py)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    print(json.dumps(line), flush=True)


# --------------------------------------------------------------------------------------------------
# our arm
# --------------------------------------------------------------------------------------------------
class ClockSampler(threading.Thread):
    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()

    def run(self):
        try:
            import pynvml as nv
            nv.nvmlInit()
            h = nv.nvmlDeviceGetHandleByIndex(self.index)
            self.max_mhz = nv.nvmlDeviceGetMaxClockInfo(h, nv.NVML_CLOCK_SM)
            names = {
                getattr(nv, "nvmlClocksEventReasonHwSlowdown", 0x8): "hw_slowdown",
                getattr(nv, "nvmlClocksEventReasonHwThermalSlowdown", 0x40): "hw_thermal_slowdown",
                getattr(nv, "nvmlClocksEventReasonSwThermalSlowdown", 0x20): "sw_thermal_slowdown",
                getattr(nv, "nvmlClocksEventReasonSwPowerCap", 0x4): "sw_power_cap",
            }
            get_reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or \
                getattr(nv, "nvmlDeviceGetCurrentClocksThrottleReasons")
            while not self._stop_evt.is_set():
                self.samples.append(nv.nvmlDeviceGetClockInfo(h, nv.NVML_CLOCK_SM))
                r = get_reasons(h)
                for bit, name in names.items():
                    if r & bit:
                        self.reasons.add(name)
                time.sleep(0.1)
        except Exception as e:  # clocks are evidence, not a dependency
            self.reasons.add(f"sampler_error:{type(e).__name__}")

    def stop(self):
        self._stop_evt.set()
        self.join(timeout=2)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons)}


class KernelTimer:
    """CUDA-event timing of chosen C-ABI entry points inside the timed region (on the launching stream)."""

    def __init__(self, names):
        self.names = set(names)
        self.records = {n: [] for n in names}
        self.enabled = False

    def install(self):
        from iit_b200 import ops
        inner = ops.call

        def timed_call(name, *a):
            if self.enabled and name in self.names:
                s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
                s.record()
                inner(name, *a)
                e.record()
                self.records[name].append((s, e, hbm_bytes(name, a) if name in HBM_ENTRY_POINTS else conv_flops(a)))
            else:
                inner(name, *a)

        ops.call = timed_call

    def summary(self):
        out = {}
        for n, recs in self.records.items():
            if recs:
                ms = [s.elapsed_time(e) for s, e, _ in recs]
                fl = [f for _, _, f in recs]
                out[n] = {"launches": len(recs), "total_ms": float(sum(ms)), "avg_ms": float(np.mean(ms))}
                if n in HBM_ENTRY_POINTS:
                    out[n]["gbs"] = float(sum(fl) / (sum(ms) * 1e-3) / 1e9)
                else:
                    out[n]["tflops"] = float(sum(fl) / (sum(ms) * 1e-3) / 1e12)
        return out


def _es(dtype_code):
    return 2 if dtype_code == 1 else 4      # IIT_BF16 = 1, IIT_F32 = 0


def hbm_bytes(name, a):
    """Algorithmic HBM bytes of one launch of an HBM-bound entry point, from its arguments (DESIGN.md section 3:
    every tensor the arithmetic needs is read once and every result written once)."""
    def ptr(v):
        return v is not None and getattr(v, "value", v) not in (None, 0)
    # a ReLU mask (one bit per element) stands in for y in the backward: 1/8 byte per element instead of _es
    if name == "iit_bn_act_fwd":            # x, residual, y, dtype, rows, c, ..., relu, relu_mask, stream
        return a[4] * a[5] * (_es(a[3]) * (2 + ptr(a[1])) + (0.125 if ptr(a[-2]) else 0))
    if name == "iit_bn_act_bwd_reduce":     # dy, y, x, dtype, rows, c, mean, invstd, relu, dgamma, dbeta, relu_mask, stream
        return a[4] * a[5] * (_es(a[3]) * 2 + ((0.125 if ptr(a[-2]) else _es(a[3])) if a[8] else 0))
    if name == "iit_bn_act_bwd_apply":      # dy, y, x, dx, dres, dtype, rows, c, ..., relu, relu_mask, stream
        return a[6] * a[7] * (_es(a[5]) * (3 + ptr(a[4])) + ((0.125 if ptr(a[-2]) else _es(a[5])) if a[-3] else 0))
    if name == "iit_bn_relu_maxpool_fwd":   # x, y, idx, xwin, dtype, n, h, w, c: read x, write y (+ xwin) (1/4 each) + idx (1 B)
        n, h, w, c = a[5:9]
        pooled = n * ((h + 1) // 2) * ((w + 1) // 2) * c
        return n * h * w * c * _es(a[4]) + pooled * (_es(a[4]) * (1 + ptr(a[3])) + 1)
    if name == "iit_bn_relu_maxpool_bwd":   # dy, idx, x, y, xwin, dx, dtype, n, h, w, c: sums read dy, y, xwin; dx reads x, dy, idx
        n, h, w, c = a[7:11]
        pooled = n * ((h + 1) // 2) * ((w + 1) // 2) * c
        return pooled * (4 * _es(a[6]) + 1) + 2 * n * h * w * c * _es(a[6])
    if name == "iit_patch_fwd":             # dst, src, box: read the box from src, write it to dst
        b = a[2]._obj
        n = b.elem_bytes
        for d in range(b.rank):
            n *= b.extent[d]
        return 2 * n
    return 0


HBM_ENTRY_POINTS = ["iit_bn_act_fwd", "iit_bn_act_bwd_reduce", "iit_bn_act_bwd_apply", "iit_bn_relu_maxpool_fwd",
                    "iit_bn_relu_maxpool_bwd", "iit_patch_fwd"]


def conv_flops(call_args):
    """Algorithmic FLOPs (2*MAC) of one convolution launch from its iit_conv_t descriptor."""
    from iit_b200._lib import Conv
    for a in call_args:
        d = getattr(a, "_obj", None)
        if isinstance(d, Conv):
            return 2.0 * d.n * d.p * d.q * d.k * d.r * d.s * d.c
    return 0.0


def _timed_region(dp, dev, fn, steps):
    dp.barrier()
    torch.cuda.synchronize()
    s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    s.record()
    for _ in range(steps):
        out = fn()
    e.record()
    torch.cuda.synchronize()
    dp.barrier()
    return dp.max_over_ranks(s.elapsed_time(e) * 1e-3, dev), out


def run_ioi_extra(rank, world, dev, dp, batch=1024, steps=3, warmup=2):
    """BASELINE.json configs[2]/[3]: IOI strict-IIT train step (small HookedTransformer LL, head-level interventions),
    data-parallel with gradient all-reduce when world > 1. bf16 activations on the tensor-core GEMMs, integer labels,
    last-position unembed (0.94 GFLOP / pair)."""
    from iit_b200.model_pairs import IOI_ModelPair
    from iit_b200.tasks.ioi import IOI_HL, corr, make_ll_model
    from oracle import restate as R
    torch.manual_seed(0)
    ll = make_ll_model(device=dev, dtype=torch.bfloat16)
    (bx, by, bv), names = R.synthetic_ioi(batch, 21 + 2 * rank, one_hot=False)
    (sx, sy, sv), _ = R.synthetic_ioi(batch, 22 + 2 * rank, one_hot=False)
    hl = IOI_HL(50257, names).to(dev)
    pair = IOI_ModelPair(hl, ll, corr, training_args={"lr": 1e-3, "batch_size": batch, "lr_scheduler": None, "cuda_graphs": True})
    if world > 1:
        pair.dp = dp
    opt = pair.make_optimizer()
    ll.train()
    base = tuple(t.to(dev) for t in (bx, by, bv))
    src = tuple(t.to(dev) for t in (sx, sy, sv))
    pair.prepare_graphs(base, src, pair.loss_fn, opt)   # 4 HL nodes + 8 strict nodes + behaviour, captured before timing
    for _ in range(warmup):
        pair.run_train_step(base, src, pair.loss_fn, opt)
    sec, m = _timed_region(dp, dev, lambda: pair.run_train_step(base, src, pair.loss_fn, opt), steps)
    return {"metric": "ioi_iit_train_pairs_per_sec", "value": batch * world * steps / sec, "unit": UNIT, "ms_per_step": 1e3 * sec / steps,
            "config": {"workload": f"IOI strict-IIT train step (train_ioi.py), HookedTransformer 6L/4H/d64/d_mlp 3072/vocab 50257, "
                                   f"batch {batch} prompts/GPU x 16 tokens, bf16 activations (fp32 master weights), integer labels, "
                                   f"last-position unembed, CUDA-graph replay per (term, node)",
                       "baseline_config": "configs[2]" if world == 1 else "configs[3]", "parallelism": f"dp{world}"},
            "steps": steps, "warmup": warmup, "dtype": "bf16", "last_step_metrics": m}


def run_pvr84_extra(rank, world, dev, dp, batch=2048, steps=3, warmup=2):
    """The headline step at the reference's DEFAULT image size (pad 7: 84x84, task_loader.py:10,43); the contract size is 56.
    Odd spatial sizes downstream (21, 11, 6, 3); 11 * F84 = 6.51 GFLOP per pair."""
    from iit_b200.model_pairs import StrictIITModelPair
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    from oracle import restate as R
    ll, hl, corr = get_alignment("mnist_pvr", {"input_shape": (1, 3, 84, 84)}, device=dev, dtype=torch.bfloat16)
    ll.load_state_dict(R.resnet18_state_dict(0))
    pair = StrictIITModelPair(hl, ll, corr, training_args={"lr": 1e-3, "cuda_graphs": True})
    if world > 1:
        pair.dp = dp
    opt = pair.make_optimizer()
    ll.train()

    def gen(seed):
        parts = [R.synthetic_pvr(min(512, batch - i), seed * 1000 + i, size=84) for i in range(0, batch, 512)]
        return tuple(torch.cat([p[j] for p in parts]).to(dev) for j in range(3))
    base, src = gen(51 + 2 * rank), gen(52 + 2 * rank)
    pair.prepare_graphs(base, src, pair.loss_fn, opt)
    for _ in range(warmup):
        pair.run_train_step(base, src, pair.loss_fn, opt)
    sec, m = _timed_region(dp, dev, lambda: pair.run_train_step(base, src, pair.loss_fn, opt), steps)
    return {"metric": "iit_train_pairs_per_sec", "value": batch * world * steps / sec, "unit": UNIT, "ms_per_step": 1e3 * sec / steps,
            "step_tflops": 11 * 592_125_440 * batch * world * steps / sec / 1e12,
            "config": {"workload": f"MNIST-PVR IIT+behaviour+strict train step at 84x84 (the reference's default pad), batch {batch} "
                                   f"pairs/GPU, bf16, CUDA-graph replay", "image": "3x84x84", "parallelism": f"dp{world}"},
            "steps": steps, "warmup": warmup, "dtype": "bf16", "last_step_metrics": m}


def run_sweep_extra(rank, world, dev, dp, batch=4096, batches_per_rank=2):
    """BASELINE.json configs[4]: the causality sweep of eval_causality.py -- 17 hook points x 4 HL nodes, counterfactuals
    generated on the device, batches sharded over ranks, counts gathered at the end. One "pair" = one (base,
    counterfactual) interchange evaluation (two LL forwards, the source one truncated at the hook point)."""
    from iit_b200 import eval_causality
    from iit_b200.tasks.mnist_pvr.counterfactual import DigitBank
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    from oracle import restate as R
    ll, _, _ = get_alignment("mnist_pvr", {"input_shape": (1, 3, 56, 56)}, device=dev, dtype=torch.bfloat16)
    ll.load_state_dict(R.resnet18_state_dict(0))
    g = torch.Generator().manual_seed(5)
    bank = DigitBank(torch.rand(2000, 1, 28, 28, generator=g).to(dev), torch.randint(0, 10, (2000,), generator=g).to(dev))
    batches = [tuple(t.to(dev) for t in R.synthetic_pvr(batch, 300 + 10 * rank + i)) for i in range(batches_per_rank)]
    args = {"batch_size": batch, "count_pairs": True}
    eval_causality.evaluate_model_on_ablations(ll, "mnist_pvr", batches[:1], args, bank, dp=dp)      # warm-up
    sec, stats = _timed_region(dp, dev, lambda: eval_causality.evaluate_model_on_ablations(ll, "mnist_pvr", batches, args, bank,
                                                                                           dp=dp), 1)
    pairs = stats.pop("_pairs") * world
    return {"metric": "iia_eval_pairs_per_sec", "value": pairs / sec, "unit": UNIT, "seconds": sec, "pairs_evaluated": pairs,
            "config": {"workload": f"causality sweep (eval_causality.py): 17 conv hook points x 4 HL nodes x {batches_per_rank} "
                                   f"batches of {batch} pairs per GPU, ResNet-18 LL bf16 eval mode, device-side counterfactuals",
                       "baseline_config": "configs[4]", "parallelism": f"dp{world}"},
            "dtype": "bf16", "mean_stat": float(np.mean([v for d in stats.values() for v in d.values()]))}


def run_ours(args, rank, world, local):
    from iit_b200 import _lib, dist as idist, ops as ops_mod
    from iit_b200.model_pairs import StrictIITModelPair
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    from oracle import restate as R  # synthetic pair generators only (shared with the CPU arm)

    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dtype = torch.bfloat16 if args.dtype == "bf16" else torch.float32
    B = args.batch
    ll, hl, corr = get_alignment("mnist_pvr", {"input_shape": (1, 3, 56, 56)}, device=dev, dtype=dtype)
    ll.load_state_dict(R.resnet18_state_dict(0))
    pair = StrictIITModelPair(hl, ll, corr, training_args={"lr": 1e-3, "cuda_graphs": bool(args.graphs)})
    dp = idist.DataParallel()
    if world > 1:
        pair.dp = dp
    opt = pair.make_optimizer()
    ll.train()
    # synthetic pairs: generated in chunks on the host, one shard per rank (different seeds per rank)
    def gen(seed):
        parts = [R.synthetic_pvr(min(512, B - i), seed * 1000 + i) for i in range(0, B, 512)]
        return tuple(torch.cat([p[j] for p in parts]).pin_memory() for j in range(3))
    base_h, src_h = gen(1 + 2 * rank), gen(2 + 2 * rank)
    base_d = tuple(t.to(dev) for t in base_h)
    src_d = tuple(t.to(dev) for t in src_h)
    h2d = 2 * sum(t.numel() * t.element_size() for t in base_h)

    def step_resident():
        return pair.run_train_step(base_d, src_d, pair.loss_fn, opt)

    # end to end: every step uploads its batch from pinned host memory (H2D inside the timed region) and reads its losses
    # back. The upload of step i+1 runs on a copy stream while step i computes, into the other of two device buffers --
    # what a prefetching loader does; the first upload of the timed region is waited for, so all K uploads are inside it.
    copy_stream = torch.cuda.Stream()
    slots = [tuple(torch.empty_like(t, device=dev) for t in (*base_h, *src_h)) for _ in range(2)]
    ready = [torch.cuda.Event(), torch.cuda.Event()]
    consumed = [torch.cuda.Event(), torch.cuda.Event()]
    e2e_state = {"i": 0, "primed": False}

    def upload(slot):
        with torch.cuda.stream(copy_stream):
            copy_stream.wait_event(consumed[slot])          # the step that last read this slot has finished
            for dst, srch in zip(slots[slot], (*base_h, *src_h)):
                dst.copy_(srch, non_blocking=True)
            ready[slot].record(copy_stream)

    def step_e2e():
        i = e2e_state["i"]
        slot = i & 1
        if not e2e_state["primed"]:
            upload(slot)
            e2e_state["primed"] = True
        torch.cuda.current_stream().wait_event(ready[slot])
        upload(slot ^ 1)                                     # next step's batch, overlapped with this step's kernels
        bufs = slots[slot]
        out = pair.run_train_step(bufs[:3], bufs[3:], pair.loss_fn, opt)  # returns host floats: the D2H read of the losses
        consumed[slot].record(torch.cuda.current_stream())
        e2e_state["i"] = i + 1
        return out

    if args.graphs:  # capture all 21 (term, node) variants before anything is timed; leaves state untouched
        pair.prepare_graphs(base_d, src_d, pair.loss_fn, opt)
    timer = KernelTimer(["iit_conv2d_fprop", "iit_conv2d_dgrad", "iit_conv2d_wgrad", "iit_conv2d_fprop_tc",
                         "iit_conv2d_dgrad_tc", "iit_conv2d_wgrad_tc", "iit_conv2d_wgrad_tc_accumulate", *HBM_ENTRY_POINTS])
    timer.install()

    def timed(fn, steps):
        dp.barrier()
        torch.cuda.synchronize()
        s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        s.record()
        for _ in range(steps):
            m = fn()
        e.record()
        torch.cuda.synchronize()
        dp.barrier()
        return dp.max_over_ranks(s.elapsed_time(e) * 1e-3, dev), m

    for _ in range(args.warmup):
        step_resident()
    if args.profile_forward:   # for `ncu --profile-from-start off`: one LL forward = exactly the fprop launches of a step
        torch.cuda.synchronize()
        torch.cuda.profiler.start()
        with torch.no_grad():
            ll(base_d[0])
        torch.cuda.synchronize()
        torch.cuda.profiler.stop()
        return
    sampler = ClockSampler(local)
    sampler.start()
    launches0 = _lib.launch_count()
    timer.enabled = not args.graphs
    if args.profile_range:
        torch.cuda.profiler.start()   # ncu --profile-from-start off: only the timed region is profiled
    sec, metrics = timed(step_resident, args.steps)
    if args.profile_range:
        torch.cuda.profiler.stop()
    timer.enabled = False
    launches = _lib.launch_count() - launches0
    clocks = sampler.stop()
    step_e2e()  # warm the copy path
    torch.cuda.synchronize()
    e2e_state["primed"] = False   # the timed region starts with its own first upload (K uploads + 1 prefetch inside it)
    sec_e2e, _ = timed(step_e2e, args.steps)
    sec_eager = None
    if args.graphs:
        # Graph replays cannot be bracketed kernel by kernel: the per-kernel CUDA-event timings (roofline, kernels)
        # come from an eager pass of the same steps, whose launches also give the per-step launch count.
        pair.training_args["cuda_graphs"] = False
        step_resident()
        launches0 = _lib.launch_count()
        timer.enabled = True
        sec_eager, _ = timed(step_resident, args.steps)
        timer.enabled = False
        launches = _lib.launch_count() - launches0
        pair.training_args["cuda_graphs"] = True

    extras = {}
    if args.extras:   # the other two metrics BASELINE.json names, measured briefly on every rank of the same job
        del base_d, src_d, base_h, src_h
        pair._graphs = {}
        pair._static = None
        ops_mod._batch_cache.clear()
        torch.cuda.empty_cache()
        for name, fn in (("iia_eval_sweep", run_sweep_extra), ("ioi_train", run_ioi_extra), ("pvr_train_84px", run_pvr84_extra)):
            try:
                extras[name] = fn(rank, world, dev, dp)
            except Exception as e:  # the headline line must survive a failure here; the failure is reported, not hidden
                extras[name] = {"error": f"{type(e).__name__}: {e}"}
    if rank != 0:
        return
    pk = peaks()
    ks = timer.summary()
    roof = None
    for n, k in ks.items():   # fraction of the measured roofline per entry point: HBM copy GB/s or sustained bf16 TFLOP/s
        k["frac_of_measured_peak"] = k["gbs"] / pk["hbm_gbs"] if "gbs" in k else k["tflops"] / pk["tflops"]
    if ks:
        top = max((n for n in ks if "tflops" in ks[n]), key=lambda n: ks[n]["total_ms"])
        k = ks[top]
        roof = {"kernel": top, "bound": "tensor", "achieved": k["tflops"], "peak": pk["tflops"], "unit": "TFLOP/s",
                "frac": k["tflops"] / pk["tflops"], "traffic": ncu_traffic(top), "peak_source": pk["src"] + " bf16 sustained",
                "share_of_step": k["total_ms"] / ((sec_eager or sec) * 1e3), "launches": k["launches"], "avg_ms": k["avg_ms"],
                "timed_in": "eager pass of the same steps (CUDA events around each launch)" if sec_eager else "timed region"}
    value = B * world * args.steps / sec
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * sec / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": args.dtype, "data": "synthetic", "config": workload_config(args, world),
        "e2e": {"value": B * world * args.steps / sec_e2e, "unit": UNIT, "h2d_bytes_per_step": h2d,
                "d2h_bytes_per_step": 3 * 4},
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "kernels": ks,
        "step_tflops": 11 * F56 * B * args.steps / sec / 1e12,
        "cuda_graphs": bool(args.graphs), "ms_per_step_eager": (1e3 * sec_eager / args.steps) if sec_eager else None,
        "last_step_metrics": metrics,
    }
    if extras:
        line["other_workloads"] = extras
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline()
    print(json.dumps(line), flush=True)


def main():
    if os.environ.get("NCCL_DEBUG", "VERSION").upper() in ("VERSION", "WARN"):
        os.environ.pop("NCCL_DEBUG", None)  # NCCL writes its version banner to stdout: keep stdout to the one JSON line
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--batch", type=int, default=4096)
    ap.add_argument("--dtype", default="bf16", choices=["bf16", "f32"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--extras", type=int, default=1, help="1: also time the IIA sweep and the IOI train step (other_workloads)")
    ap.add_argument("--profile-forward", action="store_true", help="profile one no-grad LL forward after warm-up, print nothing")
    ap.add_argument("--profile-range", action="store_true", help="cudaProfilerStart/Stop around the timed region")
    ap.add_argument("--graphs", type=int, default=1, help="1: replay each (loss term, node) variant from a CUDA graph")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if args.impl == "reference":
        run_reference(args, rank, world)
        return
    from iit_b200 import dist as idist
    rank, world, local = idist.init_from_env()
    try:
        run_ours(args, rank, world, local)
    finally:
        if world > 1:
            import torch.distributed as dist
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
