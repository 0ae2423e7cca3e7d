"""bench.py -- IIT train-step throughput (pairs/s) on the configuration BASELINE.json quotes:
MNIST-PVR, ResNet-18 LL, IIT + behaviour + strict-IIT losses, bf16 activations, batch 4096 pairs
per GPU (configs[1]); one "step" = one StrictIITModelPair.run_train_step over one batch of
synthetic (base, source) pairs = 5 LL forwards, 3 backwards, 3 clip+Adam updates.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--batch B]
                    [--dtype bf16|f32]

N>1 is launched by torchrun (one rank per GPU, NCCL gradient all-reduce, weak scaling).
Prints ONE JSON line on rank 0. `other_workloads` carries the two other metrics BASELINE.json names -- the IOI
strict-IIT train step (configs[2]/[3]) and the IIA causality sweep over the 12-node leaky correspondence (configs[4]) --
each with its own roofline, cpu_baseline and e2e.
`--impl reference` times the oracle's CPU restatement of the reference step (oracle/restate.py; the Python reference
itself needs /root/reference and transformer_lens, neither of which exists on the GPU box) on the host cores, at the
workload's own batch when that fits the time budget (the batch it ran is what its `config` says).
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


def ncu_traffic(*keys):
    """Average DRAM bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum) from the committed `ncu --set full`
    captures (profiles/ncu_traffic.json, keyed by kernel name or entry point); None if not captured."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    if not os.path.exists(path):
        return None
    table = json.load(open(path))
    for k in keys:
        ent = table.get(k)
        if ent is not None:
            return ent.get("dram_bytes_per_launch")
    return None


def workload_config(args, world, batch=None):
    batch = args.batch if batch is None else batch
    return {
        "workload": f"MNIST-PVR IIT+behaviour+strict train step, ResNet-18 LL, synthetic 56x56 4-digit grids, "
                    f"batch {batch} pairs/GPU ({'bf16' if args.dtype == 'bf16' else 'f32'} activations, "
                    f"fp32 master weights/accumulation)",
        "baseline_config": "configs[1]",
        "batch_pairs_per_gpu": batch,
        "global_batch_pairs": batch * world,
        "image": "3x56x56",
        "hook_point": HOOK,
        "mode": "q",
        "losses": ["iit", "strict", "behaviour"],
        "optimizer": "Adam lr 1e-3, clip_grad_norm 1.0, one update per loss term",
        "parallelism": f"dp{world}",
        "l2": "no flush needed: per-step activation working set is several GB (>> 126 MB L2)",
    }


# --------------------------------------------------------------------------------------------------
# reference arm / cpu baselines: the oracle's CPU restatement of the same steps (kind "port")
# --------------------------------------------------------------------------------------------------
def gen_pvr(batch, seed, size=56):
    from oracle import restate as R
    parts = [R.synthetic_pvr(min(512, batch - i), seed * 1000 + i, size=size) for i in range(0, batch, 512)]
    return tuple(torch.cat([p[j] for p in parts]) for j in range(3))


def cpu_pvr_step_factory(batch):
    from oracle import restate as R
    oll = R.ResNet18Oracle(R.resnet18_state_dict(0))
    strict = [(n, None) for n in R.ResNet18Oracle.conv_hook_points() if n != HOOK]
    pair = R.OraclePair(oll, "pvr", R.pvr_corr("q", HOOK, 256, 4), lr=1e-3, weights=(1.0, 1.0, 1.0), strict_nodes=strict)
    base, src = gen_pvr(batch, 11), gen_pvr(batch, 12)
    return lambda: pair.train_step(base, src, strict=True)


def cpu_ioi_step_factory(batch):
    """The reference's IOI strict step as written: one-hot float labels [B,16,50257], dense HL logits, all-position unembed."""
    from oracle import restate as R
    oll = R.TransformerOracle(R.transformer_state_dict(0), 6, 4, 16)
    corr = R.ioi_corr()
    strict = R.nodes_not_in_circuit(6, 4, corr, {"attn": "attn.hook_z", "mlp": "mlp.hook_post"})
    base, names = R.synthetic_ioi(batch, 21, one_hot=True)
    src, _ = R.synthetic_ioi(batch, 22, one_hot=True)
    pair = R.OraclePair(oll, "ioi", corr, lr=1e-3, strict_nodes=strict, names=names, d_vocab=50257)
    return lambda: pair.train_step(base, src, strict=True)


def cpu_sweep_factory(batch, hook_points):
    """eval_causality.py:14-97 restated: per (hook point, leaky HL node) one counterfactual batch, the interchange (two LL
    forwards in eval mode, the reference runs both in full) and the masked accuracy."""
    from oracle import restate as R
    oll = R.ResNet18Oracle(R.resnet18_state_dict(0)).eval()
    base = gen_pvr(batch, 300)
    shapes = {}
    with torch.no_grad():
        oll(base[0][:2], cache=shapes)

    def sweep():
        n = 0
        with torch.no_grad():
            for hp in hook_points:
                corr = R.pvr_leaky_corr(hp, shapes[hp].shape[2])
                for i, (node, ll_nodes) in enumerate(corr.items()):
                    src = R.synthetic_pvr_counterfactual(base, node, 500 + i)
                    hl = R.pvr_hl_intervened(base[2], src[2], node)
                    out, _ = R.ll_intervention(oll, base[0], src[0], ll_nodes)
                    R.causality_stat(out, hl, base[1], src[1])
                    n += batch
        return n

    return sweep


def time_cpu(step, units_per_step, budget_s, min_steps=3, max_steps=50, threads=None):
    """Median-of-steps throughput of a CPU step on `threads` host threads (all cores when None)."""
    cores = os.cpu_count() or 1
    torch.set_num_threads(threads or cores)
    step()  # warm-up
    times = []
    t_all = time.perf_counter()
    while len(times) < max_steps and (len(times) < min_steps or time.perf_counter() - t_all < budget_s):
        t0 = time.perf_counter()
        step()
        times.append(time.perf_counter() - t0)
    used = torch.get_num_threads()
    torch.set_num_threads(cores)
    med = float(np.median(times))
    return {"value": units_per_step / med, "cores": used, "steps": len(times), "median_s_per_step": med,
            "seconds": time.perf_counter() - t_all}


def cpu_baseline_block(step_factory, batch, budget_s, what, unit=UNIT, one_thread_batch=None):
    r = time_cpu(step_factory(batch), batch, budget_s)
    out = {"value": r["value"], "unit": unit, "cores": r["cores"], "kind": "port",
           "sample": f"median of {r['steps']} steps of {batch} pairs ({what}, fp32, torch-CPU eager restatement oracle/restate.py), "
                     f"{r['seconds']:.1f} s", "host_cpus": os.cpu_count()}
    if one_thread_batch:
        r1 = time_cpu(step_factory(one_thread_batch), one_thread_batch, budget_s / 2, min_steps=2, max_steps=10, threads=1)
        out["one_thread"] = {"value": r1["value"], "unit": unit, "cores": 1,
                             "sample": f"median of {r1['steps']} steps of {one_thread_batch} pairs, {r1['seconds']:.1f} s"}
    return out


def run_reference(args, rank, world):
    if rank != 0:
        return
    cores = os.cpu_count() or 1
    torch.set_num_threads(cores)
    total = max(args.steps + args.warmup, 1)
    # the workload's own batch when the whole run then stays within ~5 minutes, else the largest multiple of 64 that does;
    # `config` states the batch that ran
    probe = cpu_pvr_step_factory(64)
    probe()
    t0 = time.perf_counter()
    probe()
    per_pair = (time.perf_counter() - t0) / 64
    fit = int(300.0 / (total * per_pair * 0.8))      # large batches run ~20 % faster per pair than the 64-pair probe
    batch = args.batch if fit >= args.batch else max(64, fit // 64 * 64)
    step = cpu_pvr_step_factory(batch)
    for _ in range(args.warmup):
        step()
    times = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        step()
        times.append(time.perf_counter() - t0)
    dt = float(sum(times))
    value = batch * args.steps / dt
    # our arm's config, verbatim, when the workload's batch ran; otherwise the same with the batch that did run
    cfg = workload_config(args, world, batch)
    sample = (f"{args.steps} steps of {batch} pairs each on {torch.get_num_threads()} host threads (oracle/restate.py: the CPU "
              f"restatement of strict_iit_model_pair.py:36-91; the workload's batch is {args.batch}); median step "
              f"{1e3 * float(np.median(times)):.0f} ms = {batch / float(np.median(times)):.1f} pairs/s")
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": 1e3 * dt / max(args.steps, 1), "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "f32", "data": "synthetic", "config": cfg,
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": torch.get_num_threads(), "kind": "port", "sample": sample,
                         "host_cpus": cores},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    if not args.no_one_thread:
        r1 = time_cpu(cpu_pvr_step_factory(64), 64, 15.0, min_steps=2, max_steps=8, threads=1)
        line["cpu_baseline"]["one_thread"] = {"value": r1["value"], "unit": UNIT, "cores": 1,
                                              "sample": f"median of {r1['steps']} steps of 64 pairs, {r1['seconds']:.1f} s"}
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


def _es(dtype_code):
    return 2 if dtype_code == 1 else 4      # IIT_BF16 = 1, IIT_F32 = 0


def _ptr(v):
    return v is not None and getattr(v, "value", v) not in (None, 0)


def _conv_desc(call_args):
    from iit_b200._lib import Conv
    for a in call_args:
        d = getattr(a, "_obj", None)
        if isinstance(d, Conv):
            return d
    return None


def conv_flops(a):
    """Algorithmic FLOPs (2*MAC) of one convolution launch from its iit_conv_t descriptor."""
    d = _conv_desc(a)
    return 0.0 if d is None else 2.0 * d.n * d.p * d.q * d.k * d.r * d.s * d.c


def conv_bytes(name, a):
    """Algorithmic HBM bytes of one GEMM / convolution launch (bf16 operands): every operand read once, the result written once."""
    d = _conv_desc(a)
    if d is None:
        return 0.0
    x, y, w = d.n * d.h * d.w * d.c, d.n * d.p * d.q * d.k, d.k * d.r * d.s * d.c
    if "wgrad" in name:
        return 2.0 * (x + y) + 4.0 * w
    return 2.0 * (x + y + w)


def layer_key(d):
    if d.h == 1 and d.w == 1 and d.r == 1:
        return f"gemm [{d.n} x {d.c}] x [{d.c} x {d.k}]"
    return f"conv {d.r}x{d.s}/{d.stride} {d.c}->{d.k} @{d.h}x{d.w} n={d.n}"


def hbm_bytes(name, a):
    """Algorithmic HBM bytes of one launch of an HBM-bound entry point, from its arguments (DESIGN.md section 3:
    every tensor the arithmetic needs is read once and every result written once)."""
    # a ReLU mask (one bit per element) stands in for y in the backward: 1/8 byte per element instead of _es
    if name == "iit_bn_act_fwd":            # x, residual, y, dtype, rows, c, ..., relu, relu_mask, stream
        return a[4] * a[5] * (_es(a[3]) * (2 + _ptr(a[1])) + (0.125 if _ptr(a[-2]) else 0))
    if name == "iit_bn_act_bwd_reduce":     # dy, y, x, dtype, rows, c, mean, invstd, relu, dgamma, dbeta, relu_mask, stream
        return a[4] * a[5] * (_es(a[3]) * 2 + ((0.125 if _ptr(a[-2]) else _es(a[3])) if a[8] else 0))
    if name == "iit_bn_act_bwd_apply":      # dy, y, x, dx, dres, dtype, rows, c, ..., relu, relu_mask, stream
        return a[6] * a[7] * (_es(a[5]) * (3 + _ptr(a[4])) + ((0.125 if _ptr(a[-2]) else _es(a[5])) if a[-3] else 0))
    if name == "iit_bn_relu_maxpool_fwd":   # x, y, idx, xwin, dtype, n, h, w, c: read x, write y (+ xwin) (1/4 each) + idx (1 B)
        n, h, w, c = a[5:9]
        pooled = n * ((h + 1) // 2) * ((w + 1) // 2) * c
        return n * h * w * c * _es(a[4]) + pooled * (_es(a[4]) * (1 + _ptr(a[3])) + 1)
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
    # transformer pieces
    if name == "iit_bias_act_fwd":          # x, bias, resid, pre_out, out, dtype, rows, n, act
        return a[6] * a[7] * _es(a[5]) * (2 + _ptr(a[2]) + _ptr(a[3]))
    if name == "iit_gelu_bwd":              # dpost, pre, dpre, dtype, n
        return 3 * a[4] * _es(a[3])
    if name == "iit_gelu_bwd_bias_colsum":  # dpost, h, bias, dpre, colsum, rows, n (bf16)
        return 3 * a[5] * a[6] * 2
    if name == "iit_colsum":                # x, dtype, out, rows, n
        return a[3] * a[4] * _es(a[1])
    if name == "iit_attn_fwd":              # qkv, z, dtype, batch, seq, heads, d_head
        return a[3] * a[4] * a[5] * a[6] * 4 * _es(a[2])
    if name == "iit_attn_bwd":              # qkv, dz, dqkv, dtype, batch, seq, heads, d_head
        return a[4] * a[5] * a[6] * a[7] * 7 * _es(a[3])
    if name == "iit_lnpre_fwd":             # x, y, rstd, dtype, rows, d
        return a[4] * a[5] * 2 * _es(a[3])
    if name == "iit_lnpre_bwd":             # dy, y, rstd, dx, dtype, rows, d
        return a[5] * a[6] * 3 * _es(a[4])
    if name == "iit_gemm_bias_act_tc":     # x, w_oi, bias, resid, y, pre_out, rows, n_in, n_out, act: operands once, results once
        rows, n_in, n_out = a[6], a[7], a[8]
        return 2 * (rows * n_in + n_in * n_out + rows * n_out * (1 + _ptr(a[3]) + _ptr(a[5]))) + 4 * n_out * _ptr(a[2])
    if name == "iit_gemm_dgrad_gelu_tc":   # dy, w_io, pre, dpre, col_sum, col_sq, rows, n_in, n_out
        rows, n_in, n_out = a[6], a[7], a[8]
        return 2 * (rows * n_out + n_in * n_out + 2 * rows * n_in)
    if name == "iit_hl_ce_fwd":             # logits, dtype, rows, classes, row_stride, ...: one pass over the scored columns
        return a[2] * a[3] * _es(a[1])
    if name in ("iit_ce_index_fwd", "iit_ce_index_bwd"):   # logits, dtype, rows, classes, ...: one pass over the logits (+ the gradient)
        return a[2] * a[3] * _es(a[1]) * (2 if name.endswith("bwd") else 1)
    return 0


PVR_HBM = ["iit_bn_act_fwd", "iit_bn_act_bwd_reduce", "iit_bn_act_bwd_apply", "iit_bn_relu_maxpool_fwd",
           "iit_bn_relu_maxpool_bwd", "iit_patch_fwd"]
IOI_HBM = ["iit_gemm_bias_act_tc", "iit_gemm_dgrad_gelu_tc", "iit_gelu_bwd_bias_colsum", "iit_bias_act_fwd", "iit_gelu_bwd", "iit_colsum", "iit_attn_fwd", "iit_attn_bwd", "iit_lnpre_fwd", "iit_lnpre_bwd",
           "iit_hl_ce_fwd", "iit_ce_index_fwd", "iit_ce_index_bwd"]
CONV_ENTRIES = ["iit_conv2d_fprop", "iit_conv2d_dgrad", "iit_conv2d_wgrad", "iit_conv2d_fprop_tc", "iit_conv2d_dgrad_tc",
                "iit_conv2d_wgrad_tc", "iit_conv2d_wgrad_tc_accumulate", "iit_stem_conv_s2d", "iit_stem_wgrad_s2d",
                "iit_conv2d_fprop_tc_affine", "iit_conv2d_fprop_tc_patch", "iit_conv2d_dgrad_tc_accumulate", "iit_conv2d_wgrad_tc_scatter"]


def stem_flops(name, a):
    """The 7x7/2 stem through its own entry points: 2 * n * (h/2) * (w/2) * 64 * 147 algorithmic FLOPs."""
    if name == "iit_stem_conv_s2d":     # s2d, w, y, n, sh, sw, ...
        n, sh, sw = a[3], a[4], a[5]
    else:                               # iit_stem_wgrad_s2d: dy, s2d, dw, strides, n, sh, sw, ...
        n, sh, sw = a[4], a[5], a[6]
    return 2.0 * n * (sh - 3) * (sw - 3) * 64 * 147


class KernelTimer:
    """CUDA-event timing of C-ABI entry points (on the launching stream) in an eager pass. `gemm_bound`: "tensor" counts
    convolution launches in FLOPs (the ResNet), "hbm" in algorithmic bytes (the d_model-64 transformer GEMMs)."""

    def __init__(self, hbm_names, gemm_bound="tensor"):
        self.hbm, self.gemm_bound = set(hbm_names), gemm_bound
        self.records = []     # (entry, start, end, work, kind, layer, variants)
        self.enabled = False
        self._inner = None

    def install(self):
        from iit_b200 import _lib, ops
        self._inner = inner = ops.call

        def timed_call(name, *a):
            if not self.enabled or not (name in self.hbm or name in CONV_ENTRIES):
                return inner(name, *a)
            before = _lib.kernel_variant_counts() if name in CONV_ENTRIES else None
            s, e = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            s.record()
            inner(name, *a)
            e.record()
            if name in self.hbm:
                layer = None
                if name == "iit_gemm_bias_act_tc":
                    layer = (f"gemm [{a[6]} x {a[7]}] x [{a[7]} x {a[8]}] + bias{' + residual' if _ptr(a[3]) else ''}"
                             f"{' + gelu_new' if a[9] else ''}{' (pre kept)' if _ptr(a[5]) else ''}")
                elif name == "iit_gemm_dgrad_gelu_tc":
                    layer = f"gemm [{a[6]} x {a[8]}] x [{a[8]} x {a[7]}] * gelu'(pre) + column sums"
                self.records.append((name, s, e, hbm_bytes(name, a), "hbm", layer, ()))
                return
            after = _lib.kernel_variant_counts()
            variants = tuple(k for k in after if after[k] > before[k])
            d = _conv_desc(a)
            if d is None:
                work, kind, layer = stem_flops(name, a), "tensor", "stem 7x7/2 3->64 (space-to-depth)"
            elif self.gemm_bound == "hbm":
                work, kind, layer = conv_bytes(name, a), "hbm", layer_key(d)
            else:
                work, kind, layer = conv_flops(a), "tensor", layer_key(d)
            self.records.append((name, s, e, work, kind, layer, variants))

        ops.call = timed_call

    def uninstall(self):
        from iit_b200 import ops
        if self._inner is not None:
            ops.call = self._inner

    def _group(self, keyfn):
        groups = {}
        for rec in self.records:
            k = keyfn(rec)
            if k is None:
                continue
            g = groups.setdefault(k, {"launches": 0, "total_ms": 0.0, "work": 0.0, "kind": rec[4]})
            g["launches"] += 1
            g["total_ms"] += rec[1].elapsed_time(rec[2])
            g["work"] += rec[3]
        return groups

    @staticmethod
    def _finish(g, pk):
        out = {"launches": g["launches"], "total_ms": g["total_ms"], "avg_ms": g["total_ms"] / g["launches"]}
        rate = g["work"] / (g["total_ms"] * 1e-3) if g["total_ms"] > 0 else 0.0
        if g["kind"] == "hbm":
            out["gbs"] = rate / 1e9
            out["frac_of_measured_peak"] = out["gbs"] / pk["hbm_gbs"]
        else:
            out["tflops"] = rate / 1e12
            out["frac_of_measured_peak"] = out["tflops"] / pk["tflops"]
        return out

    def by_entry(self, pk):
        return {k: self._finish(g, pk) for k, g in self._group(lambda r: r[0]).items()}

    def by_layer(self, pk):
        rows = []
        for (entry, layer, variants), g in self._group(lambda r: (r[0], r[5], r[6]) if r[5] is not None else None).items():
            row = self._finish(g, pk)
            row.update({"entry": entry, "layer": layer, "kernel": " + ".join(variants) if variants else "(CUDA-core path)"})
            rows.append(row)
        return sorted(rows, key=lambda r: -r["total_ms"])

    def by_kernel(self, pk):
        """Launch groups by the kernel instantiation(s) the dispatcher picked (entries that split into several launches,
        e.g. the parity classes of a stride-2 dgrad, count under the joined name)."""
        out = {}
        for k, g in self._group(lambda r: " + ".join(r[6]) if r[6] else None).items():
            out[k] = self._finish(g, pk)
        return out


def roofline_block(timer, pk, step_ms, timed_in, prefer_kind=None):
    """The dominant kernel by summed time, named by its instantiation, with the per-layer list behind it."""
    # kernels by name: the convolution / GEMM instantiations the dispatchers picked, and the HBM entry points (each of those
    # is one kernel)
    named = {**timer.by_kernel(pk), **{k: v for k, v in timer.by_entry(pk).items() if k in timer.hbm}}
    want = "gbs" if prefer_kind == "hbm" else "tflops"
    cands = {k: v for k, v in named.items() if want in v}
    if not cands:
        return None
    top = max(cands, key=lambda n: cands[n]["total_ms"])
    k = cands[top]
    tensor = "tflops" in k
    return {"kernel": top, "bound": "tensor" if tensor else "hbm", "achieved": k["tflops"] if tensor else k["gbs"],
            "peak": pk["tflops"] if tensor else pk["hbm_gbs"], "unit": "TFLOP/s" if tensor else "GB/s",
            "frac": k["frac_of_measured_peak"], "traffic": ncu_traffic(top, top.split("<")[0]),
            "peak_source": pk["src"] + (" bf16 sustained" if tensor else " HBM copy"),
            "share_of_step": k["total_ms"] / step_ms, "launches": k["launches"], "avg_ms": k["avg_ms"], "timed_in": timed_in,
            "layers": timer.by_layer(pk)[:24]}


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


class Uploader:
    """Double-buffered upload of a step's inputs from pinned host memory on a copy stream: the upload of step i+1 overlaps
    step i's kernels (what a prefetching loader does); the first upload of a timed region is waited for, so K steps hold K
    uploads (+1 prefetch)."""

    def __init__(self, host_tensors, dev):
        self.host = [t.pin_memory() for t in host_tensors]
        self.stream = torch.cuda.Stream()
        self.slots = [tuple(torch.empty_like(t, device=dev) for t in self.host) for _ in range(2)]
        self.ready = [torch.cuda.Event(), torch.cuda.Event()]
        self.consumed = [torch.cuda.Event(), torch.cuda.Event()]
        self.i, self.primed = 0, False
        self.bytes = sum(t.numel() * t.element_size() for t in self.host)

    def _upload(self, slot):
        with torch.cuda.stream(self.stream):
            self.stream.wait_event(self.consumed[slot])          # the step that last read this slot has finished
            for dst, src in zip(self.slots[slot], self.host):
                dst.copy_(src, non_blocking=True)
            self.ready[slot].record(self.stream)

    def restart(self):
        self.primed = False

    def next(self):
        slot = self.i & 1
        if not self.primed:
            self._upload(slot)
            self.primed = True
        torch.cuda.current_stream().wait_event(self.ready[slot])
        self._upload(slot ^ 1)
        self.i += 1
        return self.slots[slot], slot

    def done(self, slot):
        self.consumed[slot].record(torch.cuda.current_stream())


def run_ioi_extra(rank, world, dev, dp, pk, batch=1024, steps=5, warmup=3):
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
    # end to end: tokens, integer labels and IO ids of both sides uploaded from pinned host memory every step, losses read back
    up = Uploader([bx, by, bv, sx, sy, sv], dev)

    def step_e2e():
        bufs, slot = up.next()
        out = pair.run_train_step(bufs[:3], bufs[3:], pair.loss_fn, opt)
        up.done(slot)
        return out

    step_e2e()
    torch.cuda.synchronize()
    up.restart()
    sec_e2e, _ = _timed_region(dp, dev, step_e2e, steps)
    # per-kernel roofline from an eager pass of the same step (graph replays cannot be bracketed launch by launch)
    timer = KernelTimer(IOI_HBM, gemm_bound="hbm")
    timer.install()
    pair.training_args["cuda_graphs"] = False
    pair.run_train_step(base, src, pair.loss_fn, opt)
    timer.enabled = True
    sec_eager, _ = _timed_region(dp, dev, lambda: pair.run_train_step(base, src, pair.loss_fn, opt), 2)
    timer.enabled = False
    timer.uninstall()
    out = {"metric": "ioi_iit_train_pairs_per_sec", "value": batch * world * steps / sec, "unit": UNIT,
           "ms_per_step": 1e3 * sec / steps, "n_gpus": world,
           "config": {"workload": f"IOI strict-IIT train step (train_ioi.py), HookedTransformer 6L/4H/d64/d_mlp 3072/vocab 50257, "
                                  f"batch {batch} prompts/GPU x 16 tokens, bf16 activations (fp32 master weights), integer labels, "
                                  f"last-position unembed, CUDA-graph replay per (term, node)",
                      "baseline_config": "configs[2]" if world == 1 else "configs[3]", "parallelism": f"dp{world}"},
           "steps": steps, "warmup": warmup, "dtype": "bf16", "last_step_metrics": m,
           "e2e": {"value": batch * world * steps / sec_e2e, "unit": UNIT, "h2d_bytes_per_step": up.bytes, "d2h_bytes_per_step": 12,
                   "note": "labels travel as int64 ids [B,16]; the reference's one-hot float labels would be 3.3 GB per side per step"},
           "ms_per_step_eager": 1e3 * sec_eager / 2}
    if rank == 0:
        ents = timer.by_entry(pk)
        # HBM-bound by construction (d_model = 64: K or N = 64 in every GEMM): the roofline of the step is
        # the sum of its kernels' algorithmic bytes over the HBM copy peak
        tot_bytes = sum(r[3] for r in timer.records)
        tot_ms = sum(r[1].elapsed_time(r[2]) for r in timer.records)
        out["roofline"] = roofline_block(timer, pk, 1e3 * sec_eager, "eager pass of the same step (CUDA events around each launch)",
                                         prefer_kind="hbm")
        out["kernels"] = ents
        out["hbm_summary"] = {"algorithmic_gb_per_step": tot_bytes / 2 / 1e9, "kernel_ms_per_step": tot_ms / 2,
                              "gbs_over_kernel_time": tot_bytes / (tot_ms * 1e-3) / 1e9 if tot_ms else None,
                              "frac_of_measured_peak": tot_bytes / (tot_ms * 1e-3) / 1e9 / pk["hbm_gbs"] if tot_ms else None,
                              "gbs_over_graph_step": tot_bytes / 2 / (sec / steps) / 1e9}
    return out


def run_pvr84_extra(rank, world, dev, dp, pk, batch=2048, steps=3, warmup=2):
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
    base = tuple(t.to(dev) for t in gen_pvr(batch, 51 + 2 * rank, size=84))
    src = tuple(t.to(dev) for t in gen_pvr(batch, 52 + 2 * rank, size=84))
    pair.prepare_graphs(base, src, pair.loss_fn, opt)
    for _ in range(warmup):
        pair.run_train_step(base, src, pair.loss_fn, opt)
    sec, m = _timed_region(dp, dev, lambda: pair.run_train_step(base, src, pair.loss_fn, opt), steps)
    return {"metric": "iit_train_pairs_per_sec", "value": batch * world * steps / sec, "unit": UNIT, "ms_per_step": 1e3 * sec / steps,
            "step_tflops": 11 * 592_125_440 * batch * world * steps / sec / 1e12,
            "config": {"workload": f"MNIST-PVR IIT+behaviour+strict train step at 84x84 (the reference's default pad), batch {batch} "
                                   f"pairs/GPU, bf16, CUDA-graph replay", "image": "3x84x84", "parallelism": f"dp{world}"},
            "steps": steps, "warmup": warmup, "dtype": "bf16", "last_step_metrics": m}


def run_sweep_extra(rank, world, dev, dp, pk, batch=4096, batches_per_rank=1):
    """BASELINE.json configs[4]: the causality sweep of eval_causality.py -- 17 conv hook points x the 12 nodes of the LEAKY
    correspondence the reference sweeps (eval_causality.py:102,138; pvr_check_leaky_hl.py:72-109), counterfactuals generated
    on the device, batches sharded over ranks, counts gathered at the end. One "pair" = one (base, counterfactual)
    interchange evaluation (two LL forwards, the source one truncated at the hook point)."""
    from iit_b200 import eval_causality
    from iit_b200.tasks.mnist_pvr.counterfactual import DigitBank
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    from oracle import restate as R
    ll, _, _ = get_alignment("mnist_pvr", {"input_shape": (1, 3, 56, 56)}, device=dev, dtype=torch.bfloat16)
    ll.load_state_dict(R.resnet18_state_dict(0))
    g = torch.Generator().manual_seed(5)
    bank = DigitBank(torch.rand(2000, 1, 28, 28, generator=g).to(dev), torch.randint(0, 10, (2000,), generator=g).to(dev))
    host_batches = [gen_pvr(batch, 300 + 10 * rank + i) for i in range(batches_per_rank)]
    batches = [tuple(t.to(dev) for t in b) for b in host_batches]
    args = {"batch_size": batch, "count_pairs": True}
    task = "pvr_leaky"
    # warm-up: one whole sweep (every hook point has its own resume point, shapes and allocator blocks; a two-point warm-up
    # left the first timed sweep 0.6 -> 1.0 s on some boxes), then the median of three timed sweeps
    eval_causality.evaluate_model_on_ablations(ll, task, batches[:1], args, bank, dp=dp)
    runs = [_timed_region(dp, dev, lambda: eval_causality.evaluate_model_on_ablations(ll, task, batches, args, bank, dp=dp), 1)
            for _ in range(3)]
    sec, stats = sorted(runs, key=lambda r: r[0])[1]
    sweep_seconds = [r[0] for r in runs]
    pairs = stats.pop("_pairs") * world
    # end to end: the base batches start in pinned host memory and are uploaded inside the timed region; the statistics
    # come back to the host at the end of the sweep (they always do: the function returns Python floats)
    pinned = [tuple(t.pin_memory() for t in b) for b in host_batches]

    def sweep_e2e():
        up = [tuple(t.to(dev, non_blocking=True) for t in b) for b in pinned]
        return eval_causality.evaluate_model_on_ablations(ll, task, up, args, bank, dp=dp)

    sec_e2e, stats2 = _timed_region(dp, dev, sweep_e2e, 1)
    stats2.pop("_pairs")
    out = {"metric": "iia_eval_pairs_per_sec", "value": pairs / sec, "unit": UNIT, "seconds": sec, "seconds_each_sweep": sweep_seconds,
           "pairs_evaluated": pairs,
           "n_gpus": world,
           "config": {"workload": f"causality sweep (eval_causality.py): 17 conv hook points x 12 leaky HL nodes x {batches_per_rank} "
                                  f"batch(es) of {batch} pairs per GPU, ResNet-18 LL bf16 eval mode, device-side counterfactuals",
                      "task": task, "baseline_config": "configs[4]", "parallelism": f"dp{world}"},
           "dtype": "bf16", "mean_stat": float(np.mean([v for d in stats.values() for v in d.values()])),
           "e2e": {"value": pairs / sec_e2e, "unit": UNIT,
                   "h2d_bytes_per_step": sum(t.numel() * t.element_size() for b in host_batches for t in b),
                   "d2h_bytes_per_step": 17 * 12 * batches_per_rank * 16}}
    if rank == 0:
        timer = KernelTimer(PVR_HBM)
        timer.install()
        timer.enabled = True
        hooks3 = ["mod.conv1.hook_point", "mod.layer2.mod.0.mod.conv1.hook_point", HOOK]
        sec_k, _ = _timed_region(idle_dp(), dev, lambda: eval_causality.evaluate_model_on_ablations(
            ll, task, batches[:1], args, bank, hook_points=hooks3), 1)
        timer.enabled = False
        timer.uninstall()
        out["roofline"] = roofline_block(timer, pk, 1e3 * sec_k, "eager sweep over 3 of the 17 hook points (CUDA events around each launch)")
        out["kernels"] = timer.by_entry(pk)
    return out


class _IdleDP:
    world, rank = 1, 0

    def barrier(self):
        pass

    def max_over_ranks(self, seconds, device):
        return float(seconds)


def idle_dp():
    return _IdleDP()


def run_ours(args, rank, world, local):
    from iit_b200 import _lib, dist as idist, ops as ops_mod
    from iit_b200.model_pairs import StrictIITModelPair
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    from oracle import restate as R  # synthetic pair generators only (shared with the CPU arm)

    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dtype = torch.bfloat16 if args.dtype == "bf16" else torch.float32
    B = args.batch
    pk = peaks()
    ll, hl, corr = get_alignment("mnist_pvr", {"input_shape": (1, 3, 56, 56)}, device=dev, dtype=dtype)
    ll.load_state_dict(R.resnet18_state_dict(0))
    pair = StrictIITModelPair(hl, ll, corr, training_args={"lr": 1e-3, "cuda_graphs": bool(args.graphs)})
    dp = idist.DataParallel()
    if world > 1:
        pair.dp = dp
    opt = pair.make_optimizer()
    ll.train()
    # synthetic pairs: generated in chunks on the host, one shard per rank (different seeds per rank)
    base_h, src_h = gen_pvr(B, 1 + 2 * rank), gen_pvr(B, 2 + 2 * rank)
    base_d = tuple(t.to(dev) for t in base_h)
    src_d = tuple(t.to(dev) for t in src_h)

    def step_resident():
        return pair.run_train_step(base_d, src_d, pair.loss_fn, opt)

    # end to end: every step uploads its batch from pinned host memory (H2D inside the timed region) and reads its losses back
    up = Uploader([*base_h, *src_h], dev)

    def step_e2e():
        bufs, slot = up.next()
        out = pair.run_train_step(bufs[:3], bufs[3:], pair.loss_fn, opt)  # returns host floats: the D2H read of the losses
        up.done(slot)
        return out

    if args.graphs:  # capture all 21 (term, node) variants before anything is timed; leaves state untouched
        pair.prepare_graphs(base_d, src_d, pair.loss_fn, opt)
    timer = KernelTimer(PVR_HBM)
    timer.install()

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
    sec, metrics = _timed_region(dp, dev, step_resident, args.steps)
    if args.profile_range:
        torch.cuda.profiler.stop()
    timer.enabled = False
    launches = _lib.launch_count() - launches0
    clocks = sampler.stop()
    step_e2e()  # warm the copy path
    torch.cuda.synchronize()
    up.restart()   # the timed region starts with its own first upload (K uploads + 1 prefetch inside it)
    sec_e2e, _ = _timed_region(dp, dev, step_e2e, args.steps)
    # the other end-to-end path: the dataset lives on the device (DeviceIITDataset); a step uploads
    # only the pair indices of its batch (2 x B int64), gathers the batch with the gather kernels and reads the losses back
    from iit_b200.utils.iit_dataset import DeviceIITDataset
    ds = DeviceIITDataset(base_d, src_d, seed=rank)
    ds.pair_table()
    perm_rng = np.random.default_rng(1234 + rank)

    def step_dataset():
        base, abl = ds.gather_batch(perm_rng.permutation(B))
        return pair.run_train_step(base, abl, pair.loss_fn, opt)

    step_dataset()
    sec_ds, _ = _timed_region(dp, dev, step_dataset, args.steps)
    # data-parallel consistency: every rank must hold the same parameters after the same updates
    chk = opt.flat_p.double().sum().reshape(1)
    if world > 1:
        import torch.distributed as tdist
        allc = [torch.empty_like(chk) for _ in range(world)]
        tdist.all_gather(allc, chk)
        dp_check = float(max(abs(float(c) - float(allc[0])) for c in allc))
    else:
        dp_check = 0.0
    sec_eager = None
    if args.graphs:
        # Graph replays cannot be bracketed kernel by kernel: the per-kernel CUDA-event timings (roofline, kernels)
        # come from an eager pass of the same steps, whose launches also give the per-step launch count.
        pair.training_args["cuda_graphs"] = False
        step_resident()
        launches0 = _lib.launch_count()
        timer.enabled = True
        sec_eager, _ = _timed_region(dp, dev, step_resident, args.steps)
        timer.enabled = False
        launches = _lib.launch_count() - launches0
        pair.training_args["cuda_graphs"] = True
    timer.uninstall()

    extras = {}
    if args.extras:   # the other two metrics BASELINE.json names, measured briefly on every rank of the same job
        del base_d, src_d
        pair._graphs = {}
        pair._static = None
        ops_mod._batch_cache.clear()
        torch.cuda.empty_cache()
        for name, fn in (("iia_eval_sweep", run_sweep_extra), ("ioi_train", run_ioi_extra), ("pvr_train_84px", run_pvr84_extra)):
            try:
                extras[name] = fn(rank, world, dev, dp, pk)
            except Exception as e:  # the headline line must survive a failure here; the failure is reported, not hidden
                extras[name] = {"error": f"{type(e).__name__}: {e}"}
            torch.cuda.empty_cache()
    if rank != 0:
        return
    step_ms = (sec_eager or sec) * 1e3 / args.steps
    timed_in = "eager pass of the same steps (CUDA events around each launch)" if sec_eager else "timed region"
    roof = roofline_block(timer, pk, step_ms * args.steps, timed_in)
    value = B * world * args.steps / sec
    line = {
        "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps, "warmup": args.warmup,
        "ms_per_step": 1e3 * sec / args.steps, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
        "dtype": args.dtype, "data": "synthetic", "config": workload_config(args, world),
        "e2e": {"value": B * world * args.steps / sec_e2e, "unit": UNIT, "h2d_bytes_per_step": up.bytes,
                "d2h_bytes_per_step": 3 * 4},
        "e2e_device_dataset": {"value": B * world * args.steps / sec_ds, "unit": UNIT, "h2d_bytes_per_step": 2 * B * 8,
                               "d2h_bytes_per_step": 3 * 4,
                               "note": "DeviceIITDataset: samples resident in HBM, per step the pair indices go up and six gather kernels build the batch"},
        "gpu_launches": int(launches), "clocks": clocks, "roofline": roof, "kernels": timer.by_entry(pk),
        "kernel_instantiations": timer.by_kernel(pk),
        "step_tflops": 11 * F56 * B * args.steps / sec / 1e12, "dp_check": dp_check,
        "cuda_graphs": bool(args.graphs), "ms_per_step_eager": (1e3 * sec_eager / args.steps) if sec_eager else None,
        "last_step_metrics": metrics,
    }
    if extras:
        line["other_workloads"] = extras
    if world == 1 and not args.no_cpu_baseline:
        line["cpu_baseline"] = cpu_baseline_block(cpu_pvr_step_factory, 256, 20.0, "same step: IIT+strict+behaviour",
                                                  one_thread_batch=64)
        if "ioi_train" in extras and "error" not in extras["ioi_train"]:
            extras["ioi_train"]["cpu_baseline"] = cpu_baseline_block(
                cpu_ioi_step_factory, 64, 10.0, "the reference's IOI strict step: one-hot labels, dense HL logits, all-position unembed")
        if "iia_eval_sweep" in extras and "error" not in extras["iia_eval_sweep"]:
            hooks3 = ["mod.conv1.hook_point", "mod.layer2.mod.0.mod.conv1.hook_point", HOOK]
            sweep = cpu_sweep_factory(128, hooks3)
            r = time_cpu(sweep, 3 * 12 * 128, 10.0, min_steps=1, max_steps=5)
            extras["iia_eval_sweep"]["cpu_baseline"] = {
                "value": r["value"], "unit": UNIT, "cores": r["cores"], "kind": "port", "host_cpus": os.cpu_count(),
                "sample": f"median of {r['steps']} sweeps of 3 hook points x 12 leaky nodes x 128 pairs (eval_causality.py:14-97 "
                          f"restated, both LL forwards in full as the reference runs them), {r['seconds']:.1f} s"}
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
    ap.add_argument("--no-one-thread", action="store_true", help="reference arm: skip the extra one-thread sample")
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
    run_ours(args, rank, world, local)
    if world > 1:
        # leave together, then end without tearing the communicator down: captured graphs still hold its collectives and
        # destroying it under them can block at exit (seen in tests/dp_grad_check.py)
        import torch.distributed as dist
        done = torch.ones(1, device=torch.device("cuda", local))
        dist.all_reduce(done)
        torch.cuda.synchronize()
        sys.stdout.flush()
        sys.stderr.flush()
        os._exit(0)


if __name__ == "__main__":
    main()
