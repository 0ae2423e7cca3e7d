"""Run under torchrun on N >= 2 GPUs of one box (tests/test_gpu_multi.py launches it; gpurun --gpus 2 for a manual run):

    python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 tests/dp_grad_check.py

SURVEY.md 8(e): data-parallel correctness on hardware. For the IOI transformer (no batch statistics, so N-GPU and 1-GPU
runs compute the same function) every rank takes its shard of one global batch, runs each loss term of the strict step with
the gradient all-reduce (two buckets, the tail one launched from inside the backward), and rank 0 compares the averaged
gradient with the gradient of the WHOLE batch on one GPU: relative L2 <= 1e-6 in fp32. Then K free-running steps: every rank
must hold bit-identical parameters (max |checksum difference| == 0). For the ResNet the same identity holds for the
AVERAGED gradient of per-rank batches (BatchNorm statistics are per rank by design): checked against the mean of the
single-GPU gradients of the two shards."""
import json
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def flat_grad(ll):
    return torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).detach().flatten().double() for p in ll.parameters()])


def say(msg):
    print(f"[rank {os.environ.get('RANK', '0')}] {msg}", file=sys.stderr, flush=True)


def main():
    import faulthandler
    faulthandler.dump_traceback_later(120, exit=True)      # a hang prints every thread's stack and ends the process
    from iit_b200 import dist as idist
    from iit_b200.model_pairs import IOI_ModelPair, StrictIITModelPair
    from iit_b200.tasks.ioi import IOI_HL, corr, make_ll_model
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    from oracle import restate as R
    rank, world, local = idist.init_from_env()
    assert world >= 2, "run under torchrun with at least 2 ranks"
    dev = torch.device("cuda", local)
    torch.cuda.set_device(dev)
    dp = idist.DataParallel()
    out = {"world": world}

    # ---- IOI: N-GPU averaged gradient == 1-GPU gradient of the global batch ------------------------------------------
    Bg = 64 * world
    (bx, by, bv), names = R.synthetic_ioi(Bg, 21, one_hot=False)
    (sx, sy, sv), _ = R.synthetic_ioi(Bg, 22, one_hot=False)
    sd = R.transformer_state_dict(0)

    def make_ioi(use_dp):
        ll = make_ll_model(device=dev, dtype=torch.float32)
        ll.load_state_dict(sd)
        pair = IOI_ModelPair(IOI_HL(50257, names).to(dev), ll, corr, training_args={"lr": 1e-3, "batch_size": Bg, "lr_scheduler": None})
        if use_dp:
            pair.dp = dp
        ll.train()
        return pair, ll, pair.make_optimizer()

    gb, gs = tuple(t.to(dev) for t in (bx, by, bv)), tuple(t.to(dev) for t in (sx, sy, sv))
    pair, ll, opt = make_ioi(True)
    lb, ls = dp.shard(gb), dp.shard(gs)
    hl_node, ll_node = pair.sample_hl_name(), pair.sample_ll_node()
    terms = {"iit": lambda p, b, s: p.get_IIT_loss_over_batch(b, s, hl_node, p.loss_fn),
             "strict": lambda p, b, s: p.get_strict_loss_over_batch(b, s, ll_node, p.loss_fn),
             "behaviour": lambda p, b, s: p.get_behaviour_loss_over_batch(b, p.loss_fn)}
    dp_grads = {}
    for name, fn in terms.items():
        opt.zero_grad()
        fn(pair, lb, ls).backward()
        dp.allreduce_grads(opt, ll)
        dp_grads[name] = flat_grad(ll)
    say("ioi: data-parallel gradients done")
    single, sll, sopt = make_ioi(False)        # every rank: the whole global batch on its own GPU, no collective
    errs = {}
    for name, fn in terms.items():
        sopt.zero_grad()
        fn(single, gb, gs).backward()
        g1 = flat_grad(sll)
        errs[name] = float((dp_grads[name] - g1).norm() / g1.norm())
    out["ioi_grad_rel_l2_vs_1gpu"] = errs
    say(f"ioi: single-GPU comparison done {errs}")
    # K free-running strict steps with graphs: identical parameters on every rank
    pair2, ll2, opt2 = make_ioi(True)
    pair2.training_args["cuda_graphs"] = True
    for _ in range(4):
        pair2.run_train_step(lb, ls, pair2.loss_fn, opt2)
    chk = opt2.flat_p.double().sum().reshape(1)
    allc = [torch.empty_like(chk) for _ in range(world)]
    dist.all_gather(allc, chk)
    out["ioi_param_checksum_max_diff"] = float(max(abs(float(c) - float(allc[0])) for c in allc))
    say("ioi: free-running graph steps done")

    # ---- ResNet: averaged gradient == mean of the shards' single-GPU gradients (BatchNorm statistics are per rank) ---
    hook = "mod.layer3.mod.1.mod.conv2.hook_point"
    Br = 32
    base = tuple(t.to(dev) for t in R.synthetic_pvr(Br, 100 + rank))
    src = tuple(t.to(dev) for t in R.synthetic_pvr(Br, 200 + rank))
    rll, rhl, rcorr = get_alignment("mnist_pvr", {"hook_point": hook, "input_shape": (1, 3, 56, 56)}, device=dev)
    rll.load_state_dict(R.resnet18_state_dict(0))
    rpair = StrictIITModelPair(rhl, rll, rcorr, training_args={"lr": 1e-3})
    rpair.dp = dp
    ropt = rpair.make_optimizer()
    rll.train()
    node = rpair.sample_hl_name()
    ropt.zero_grad()
    rpair.get_IIT_loss_over_batch(base, src, node, rpair.loss_fn).backward()
    local_g = flat_grad(rll)
    dp.allreduce_grads(ropt, rll)
    avg_g = flat_grad(rll)
    parts = [torch.empty_like(local_g) for _ in range(world)]
    dist.all_gather(parts, local_g)
    want = torch.stack(parts).mean(0)
    out["resnet_avg_grad_rel_l2_vs_mean_of_shards"] = float((avg_g - want).norm() / want.norm())
    say("resnet: averaged gradient done")
    for _ in range(3):
        rpair.run_train_step(base, src, rpair.loss_fn, ropt)
    chk = ropt.flat_p.double().sum().reshape(1)
    allc = [torch.empty_like(chk) for _ in range(world)]
    dist.all_gather(allc, chk)
    out["resnet_param_checksum_max_diff"] = float(max(abs(float(c) - float(allc[0])) for c in allc))
    if rank == 0:
        print("DP_CHECK " + json.dumps(out), flush=True)
        assert all(v <= 1e-6 for v in out["ioi_grad_rel_l2_vs_1gpu"].values()), out
        assert out["ioi_param_checksum_max_diff"] == 0.0 and out["resnet_param_checksum_max_diff"] == 0.0, out
        assert out["resnet_avg_grad_rel_l2_vs_mean_of_shards"] <= 1e-6, out
    faulthandler.cancel_dump_traceback_later()
    # leave together (a tiny all-reduce the host waits for), then end the process without tearing NCCL down: the captured
    # graphs above still hold collectives of this communicator, and destroying it under them blocked at exit
    done = torch.ones(1, device=dev)
    dist.all_reduce(done)
    torch.cuda.synchronize()
    say("done")
    sys.stdout.flush()
    sys.stderr.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
