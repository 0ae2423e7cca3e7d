"""Parity at the CONTRACT configurations of BASELINE.json, on the kernels the bench runs:

  C2  MNIST-PVR IIT + behaviour + strict, bf16, batch 4096, CUDA-graph replay, CTA-pair convolution kernel
      (strict_iit_model_pair.py:36-91 of the reference);
  C1  MNIST-PVR IIT + behaviour, fp32, batch 256 (train.py default; iit_behavior_model_pair.py:66-141);
  C3  IOI strict step, batch 256, the reference's one-hot float labels [B,16,50257] and full vocabulary
      (train_ioi.py:16-48, ioi_dataset_tl.py:375-377).

Every loss term is teacher-forced: both sides start the term from the SAME weights (synchronised from the CPU oracle,
oracle/restate.py), so the comparison is the kernels' arithmetic and not two chaotic free-running trajectories
(tests/test_gpu_pvr_parity.py explains why). Bars: fp32 losses 1e-3 relative and patched activations 1e-3 of their range
(north_star); bf16 stated separately, tightened to what was measured at batch 4096 (printed by the test; DESIGN.md
section 5 records the numbers)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import restate as R  # noqa: E402

HOOK = "mod.layer3.mod.1.mod.conv2.hook_point"
DEV = "cuda"


def dev(inp):
    return tuple(t.to(DEV) for t in inp)


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-12)


def gen_pvr(batch, seed):
    parts = [R.synthetic_pvr(min(512, batch - i), seed * 1000 + i) for i in range(0, batch, 512)]
    return tuple(torch.cat([p[j] for p in parts]) for j in range(3))


def grad_rel_l2(ll, oll):
    num = den = 0.0
    for k, p in ll.named_parameters():
        g, og = p.grad.detach().cpu().double(), oll.p[k].grad.double()
        num += float(((g - og) ** 2).sum())
        den += float((og ** 2).sum())
    return (num / den) ** 0.5


def act_err(got, want):
    """max |diff| relative to the tensor's range, and relative L2."""
    got, want = got.float().cpu(), want.float()
    rng = float(want.max() - want.min())
    return float((got - want).abs().max()) / rng, float((got - want).norm() / want.norm())


def _pvr_pair(cls, dtype, targs):
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    ll, hl, corr = get_alignment("mnist_pvr", {"mode": "q", "hook_point": HOOK, "input_shape": (1, 3, 56, 56)}, dtype=dtype)
    ll.load_state_dict(R.resnet18_state_dict(0))
    return cls(hl, ll, corr, training_args={"lr": 1e-3, **targs})


def _hook_activations(pair, oll, b, s, node_name):
    """(source activation at the hook, patched base activation at the hook) from both sides, same weights."""
    from iit_b200.model_pairs.nodes import HLNode
    ll = pair.ll_model
    ll_nodes = pair.corr[HLNode(node_name, -1)]
    seen = {}
    with torch.no_grad():
        pair.ll_source_run(s[0].to(DEV), ll_nodes)
        src_act = pair.ll_cache[HOOK].clone()
        hooks = [(n.name, pair.make_ll_ablation_hook(n)) for n in ll_nodes]
        hooks.append((HOOK, lambda t, hook: seen.__setitem__("patched", t.detach().clone())))
        ll.run_with_hooks(b[0].to(DEV), fwd_hooks=hooks)
        ocache = {}
        oll(s[0], cache=ocache, cache_filter={HOOK})
        spec = R.pvr_corr("q", HOOK, 256, 4)[node_name]
        pcache = {}
        oll(b[0], hooks={n: (lambda act, n=n, sp=sp: R.patch(act, ocache[n], sp)) for n, sp in spec}, cache=pcache,
            cache_filter={HOOK})
    return (src_act, ocache[HOOK]), (seen["patched"], pcache[HOOK])


def test_c2_batch4096_bf16_graphs_pair_kernel_teacher_forced():
    """C2 as bench.py runs it: batch 4096, bf16 activations, every loss term replayed from its CUDA graph, layer3 / layer4
    on the CTA-pair kernel and layer2 on the two-pixel-box tiles. Per term, from synchronised weights: loss and the
    gradient the graph leaves in the optimizer's flat buffer against the fp32 CPU oracle; then the source and patched
    activations at the hooked layer."""
    from iit_b200 import _lib
    from iit_b200.model_pairs import StrictIITModelPair
    B = 4096
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    pair = _pvr_pair(StrictIITModelPair, torch.bfloat16, {"cuda_graphs": True})
    ll = pair.ll_model
    opt = pair.make_optimizer()
    ll.train()
    oll = R.ResNet18Oracle(R.resnet18_state_dict(0))
    strict_names = [n for n in R.ResNet18Oracle.conv_hook_points() if n != HOOK]
    opair = R.OraclePair(oll, "pvr", R.pvr_corr("q", HOOK, 256, 4), lr=1e-3, strict_nodes=[(n, None) for n in strict_names])
    b, s = gen_pvr(B, 11), gen_pvr(B, 12)
    sb, ss = pair._static_inputs(dev(b), dev(s))
    hl_node = pair.sample_hl_name()
    assert str(hl_node) == opair.sample_hl_name()
    ll_node = pair.sample_ll_node()
    o_ll_node = opair.sample_ll_node()
    assert ll_node.name == o_ll_node[0]
    before = _lib.kernel_variant_counts()
    terms = [
        ("iit", lambda: pair._graphed_iit(sb, ss, hl_node, pair.loss_fn, opt), lambda: opair.iit_loss(b, s, str(hl_node))),
        ("strict", lambda: pair._graphed_strict(sb, ss, ll_node, pair.loss_fn, opt),
         lambda: opair.loss_fn(R.ll_intervention(oll, b[0], s[0], [o_ll_node])[0], b[1])),
        ("behaviour", lambda: pair._graphed_behavior(sb, pair.loss_fn, opt), lambda: opair.loss_fn(oll(b[0]), b[1])),
    ]
    worst_loss = worst_grad = 0.0
    for name, mine_fn, oracle_fn in terms:
        for rep in range(2):   # the first call captures (and runs) the graph, the second is a pure replay: both from synced weights
            ll.load_state_dict(oll.state_dict())
            lm = float(mine_fn().item())        # graph: loss -> zero_grad -> backward -> clip + Adam; gradient stays in flat_g
            if rep == 0:
                opair.opt.zero_grad()
                lo = oracle_fn()
                lo.backward()
            e_loss, e_grad = rel(lm, lo.item()), grad_rel_l2(ll, oll)
            print(f"C2 {name} rep {rep}: loss {lm:.6f} vs oracle {lo.item():.6f} (rel {e_loss:.2e}), grad rel-L2 {e_grad:.2e}")
            worst_loss, worst_grad = max(worst_loss, e_loss), max(worst_grad, e_grad)
            # measured at this configuration: loss 2e-5 .. 5e-5 relative, so the bf16 loss bar here is the fp32 bar, 1e-3.
            # Gradient 0.21 in relative L2 (cosine 0.93 at the stem, 0.99 at layer4): that is what ROUNDING THE FORWARD
            # ACTIVATIONS TO bf16 does to the fp32 oracle's own gradient at this random initialisation (0.35 at batch
            # 256, tests/test_bf16_rounding_model.py), not kernel error -- the fp32 kernels agree to 2e-3 (C1 below).
            assert e_loss < 1e-3, (name, rep, lm, lo.item())
            assert e_grad < 3e-1, (name, rep, e_grad)
        torch.nn.utils.clip_grad_norm_(oll.parameters(), 1.0)
        opair.opt.step()
    after = _lib.kernel_variant_counts()
    ran = {k for k in after if after[k] > before[k]}
    # the kernels bench.py's headline runs on: CTA pairs (layers 3-4 with 256-filter tiles, layer2 with 128-filter tiles),
    # shifted-window layer1 / stem, tensor-core weight gradients
    for k in ("conv3x3_tc2_kernel<256,6>", "conv3x3_tc2_kernel<128,8>", "conv_win_kernel<128,9,4>", "conv_win_kernel<32,16,4>",
              "wgrad3x3_tc_kernel", "wgrad_win_kernel", "stem_wgrad_win_kernel"):
        assert k in ran, (k, ran)
    # activations at the hooked layer, eager, same (oracle) weights, train-mode statistics
    ll.load_state_dict(oll.state_dict())
    (sa, so), (pa, po) = _hook_activations(pair, oll, b, s, str(hl_node))
    es, ep = act_err(sa, so), act_err(pa, po)
    print(f"C2 activations at {HOOK}: source max/range {es[0]:.2e} rel-L2 {es[1]:.2e}; patched max/range {ep[0]:.2e} "
          f"rel-L2 {ep[1]:.2e}; worst loss rel {worst_loss:.2e}, worst grad rel-L2 {worst_grad:.2e}")
    # measured: 2.4e-2 in relative L2 and 1.2e-2 of the range for both (13 convolution + BatchNorm layers of bf16 storage)
    assert es[1] < 4e-2 and ep[1] < 4e-2, (es, ep)
    assert es[0] < 3e-2 and ep[0] < 3e-2, (es, ep)


def test_c1_batch256_fp32_teacher_forced():
    """C1 (train.py default): batch 256, fp32, IIT + behaviour. Losses 1e-3, gradients 1e-2 relative L2, the source and
    the patched activation at mod.layer3.mod.1.mod.conv2.hook_point within 1e-3 of their range."""
    from iit_b200.model_pairs import IITBehaviorModelPair
    B = 256
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    pair = _pvr_pair(IITBehaviorModelPair, torch.float32, {})
    ll = pair.ll_model
    opt = pair.make_optimizer()
    ll.train()
    oll = R.ResNet18Oracle(R.resnet18_state_dict(0))
    opair = R.OraclePair(oll, "pvr", R.pvr_corr("q", HOOK, 256, 4), lr=1e-3)
    for step in range(2):
        b, s = R.synthetic_pvr(B, 100 + step), R.synthetic_pvr(B, 200 + step)
        bd, sd_ = dev(b), dev(s)
        hl_node = pair.sample_hl_name()
        assert str(hl_node) == opair.sample_hl_name()
        for name, mine_fn, oracle_fn in [
                ("iit", lambda: pair.get_IIT_loss_over_batch(bd, sd_, hl_node, pair.loss_fn), lambda: opair.iit_loss(b, s, str(hl_node))),
                ("behaviour", lambda: pair.get_behaviour_loss_over_batch(bd, pair.loss_fn), lambda: opair.loss_fn(oll(b[0]), b[1]))]:
            ll.load_state_dict(oll.state_dict())
            opt.zero_grad()
            lm = mine_fn()
            lm.backward()
            opair.opt.zero_grad()
            lo = oracle_fn()
            lo.backward()
            e = grad_rel_l2(ll, oll)
            print(f"C1 step {step} {name}: loss rel {rel(lm.item(), lo.item()):.2e}, grad rel-L2 {e:.2e}")
            assert rel(lm.item(), lo.item()) < 1e-3, (step, name, lm.item(), lo.item())
            assert e < 1e-2, (step, name, e)
            torch.nn.utils.clip_grad_norm_(oll.parameters(), 1.0)
            opair.opt.step()
        ll.load_state_dict(oll.state_dict())
        (sa, so), (pa, po) = _hook_activations(pair, oll, b, s, str(hl_node))
        es, ep = act_err(sa, so), act_err(pa, po)
        print(f"C1 step {step} activations: source {es}, patched {ep}")
        assert es[0] < 1e-3 and ep[0] < 1e-3, (es, ep)


@pytest.mark.parametrize("dtype,loss_bar,grad_bar", [(torch.float32, 1e-3, 1e-2), (torch.bfloat16, 2e-2, 1.5e-1)])
def test_c3_ioi_batch256_one_hot_labels_full_vocab(dtype, loss_bar, grad_bar):
    """C3: the IOI strict step at batch 256 with the label format the reference's dataset hands over -- one-hot float
    [B,16,50257] (ioi_dataset_tl.py:375-377) -- and the full 50257-token vocabulary, all three terms teacher-forced
    against the CPU oracle (which builds the dense HL logits and one-hot targets the way ioi_model_pair.py:55-69 does)."""
    from iit_b200.model_pairs import IOI_ModelPair
    from iit_b200.tasks.ioi import IOI_HL, corr, make_ll_model
    B, Vf = 256, 50257
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    sd = R.transformer_state_dict(0)
    (bx, by, bv), names = R.synthetic_ioi(B, 31, one_hot=True)
    (sx, sy, sv), _ = R.synthetic_ioi(B, 32, one_hot=True)
    assert by.shape == (B, 16, Vf) and by.dtype == torch.float32
    b, s = (bx, by, bv), (sx, sy, sv)
    ll = make_ll_model(device=DEV, dtype=dtype)
    ll.load_state_dict(sd)
    hl = IOI_HL(Vf, names).to(DEV)
    pair = IOI_ModelPair(hl, ll, corr, training_args={"lr": 1e-3, "batch_size": B, "lr_scheduler": None})
    opt = pair.make_optimizer()
    ll.train()
    oll = R.TransformerOracle(sd, 6, 4, 16)
    ocorr = R.ioi_corr()
    strict = R.nodes_not_in_circuit(6, 4, ocorr, {"attn": "attn.hook_z", "mlp": "mlp.hook_post"})
    opair = R.OraclePair(oll, "ioi", ocorr, lr=1e-3, strict_nodes=strict, names=names, d_vocab=Vf)
    bd, sd_ = dev(b), dev(s)
    hl_node = pair.sample_hl_name()
    assert str(hl_node) == opair.sample_hl_name()
    ll_node = pair.sample_ll_node()
    o_ll_node = opair.sample_ll_node()
    assert ll_node.name == o_ll_node[0]
    # the HL model's intervened last-position argmax is integer work: equal
    hl_dense = R.ioi_hl_intervened(b[0], s[0], names, Vf, str(hl_node))
    got_am = pair.hl_model.interchange_last_argmax(bd, sd_, pair._resolve_hl_node(hl_node))
    assert torch.equal(got_am.cpu(), torch.argmax(hl_dense[:, -1, :], dim=-1))
    terms = [
        ("iit", lambda: pair.get_IIT_loss_over_batch(bd, sd_, hl_node, pair.loss_fn), lambda: opair.iit_loss(b, s, str(hl_node))),
        ("strict", lambda: pair.get_strict_loss_over_batch(bd, sd_, ll_node, pair.loss_fn),
         lambda: opair.loss_fn(R.ll_intervention(oll, b[0], s[0], [o_ll_node])[0], b[1])),
        ("behaviour", lambda: pair.get_behaviour_loss_over_batch(bd, pair.loss_fn), lambda: opair.loss_fn(oll(b[0]), b[1])),
    ]
    for name, mine_fn, oracle_fn in terms:
        ll.load_state_dict({k: v.detach() for k, v in oll.p.items()})
        opt.zero_grad()
        lm = mine_fn()
        lm.backward()
        opair.opt.zero_grad()
        lo = oracle_fn()
        lo.backward()
        num = den = 0.0
        for k, p in ll.named_parameters():
            og = oll.p[k].grad
            og = torch.zeros_like(oll.p[k]).double() if og is None else og.double()
            g = p.grad.detach().cpu().double()
            num += float(((g - og) ** 2).sum())
            den += float((og ** 2).sum())
        e = (num / den) ** 0.5
        print(f"C3 {dtype} {name}: loss {lm.item():.6f} vs {lo.item():.6f} (rel {rel(lm.item(), lo.item()):.2e}), grad rel-L2 {e:.2e}")
        assert rel(lm.item(), lo.item()) < loss_bar, (name, lm.item(), lo.item())
        assert e < grad_bar, (name, e)
        torch.nn.utils.clip_grad_norm_(oll.parameters(), 1.0)
        opair.opt.step()
