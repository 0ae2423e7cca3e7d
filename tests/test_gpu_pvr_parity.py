"""End-to-end parity of the PVR model pair on the GPU against
  (1) tests/golden/pvr.npz -- outputs of the UNMODIFIED reference (oracle/make_golden.py), and
  (2) oracle/restate.py run on the CPU on the same seeded pairs and weights.
Bars (BASELINE.json north_star): HL outputs, node selection and IIA counts bit-exact; losses and
activations within 1e-3 relative in fp32. bf16 is stated separately (looser, vs the fp32 oracle)."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import restate as R  # noqa: E402

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "pvr.npz"))
HOOK = "mod.layer3.mod.1.mod.conv2.hook_point"
DEV = "cuda"
B = int(GOLD["B"])


def close_after_updates(got, want, what):
    """Parameters after several clip+Adam updates. Adam's step is ~lr*sign(g) early on, so an element whose
    gradient is at rounding level may legitimately land 2*lr apart; the bulk must still agree closely."""
    d = np.abs(got - want)
    assert d.mean() < 1e-3 and d.max() < 2e-2, f"{what}: mean |diff| {d.mean():.3e}, max {d.max():.3e}"


def dev(inp):
    return tuple(t.to(DEV) for t in inp)


def make_pair(mode="q", cls=None, dtype=torch.float32, targs=None, task="mnist_pvr", hook=HOOK):
    from iit_b200.model_pairs import IITBehaviorModelPair
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    ll, hl, corr = get_alignment(task, {"mode": mode, "hook_point": hook, "input_shape": (1, 3, 56, 56)}, dtype=dtype)
    ll.load_state_dict(R.resnet18_state_dict(0))  # after get_corr: its probe forward moves BN running stats
    cls = cls or IITBehaviorModelPair
    return cls(hl, ll, corr, training_args={"lr": 1e-3, **(targs or {})})


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-12)


@pytest.mark.parametrize("mode", ["q", "c"])
def test_fp32_matches_reference_golden(mode):
    pair = make_pair(mode)
    ll = pair.ll_model
    ll.train()
    base, src = dev(R.synthetic_pvr(B, 100)), dev(R.synthetic_pvr(B, 200))
    sd0 = {k: v.clone() for k, v in ll.state_dict().items()}
    for k in pair.corr:
        hl, llo = pair.do_intervention(base, src, k)
        assert np.array_equal(hl.cpu().numpy(), GOLD[f"{mode}/di/{k.name}/hl"])
        np.testing.assert_allclose(llo.detach().cpu().numpy(), GOLD[f"{mode}/di/{k.name}/ll"], rtol=1e-3, atol=1e-4)
        if mode == "q" and k.name == "hook_tr":
            np.testing.assert_allclose(pair.ll_cache[HOOK][:4].cpu().numpy(), GOLD["q/di/hook_tr/cache_hook"],
                                       rtol=1e-3, atol=1e-4)
    ll.load_state_dict(sd0)
    pair.rng = np.random.default_rng(0)
    opt = pair.make_optimizer()
    nodes = []
    for step in range(3):
        base, src = dev(R.synthetic_pvr(B, 100 + step)), dev(R.synthetic_pvr(B, 200 + step))
        state = pair.rng.bit_generator.state
        nodes.append(str(pair.sample_hl_name()))
        pair.rng.bit_generator.state = state
        m = pair.run_train_step(base, src, pair.loss_fn, opt)
        # The first loss is computed from the pinned weights: 1e-3 bar. Every later loss follows >= 1 clip+Adam
        # update; free-running trajectories of two fp32 implementations separate there (one ReLU mask flipped by
        # rounding changes a whole BatchNorm channel's gradient, and Adam's early step is ~lr*sign(g)), which the
        # reference's own CPU results show across host CPUs too. Free-running bars: 1e-2 after one update, a 1e-1
        # sanity bound later. The 1e-3 bar for
        # every loss term is enforced from synchronised weights in test_teacher_forced_terms_match_oracle.
        # (tests/test_oracle_golden.py shows the same: the CPU oracle on a different host CPU leaves the
        # golden trajectory by >1e-4 after the first update and by percents two steps later.)
        tol_iit, tol_beh = (1e-3, 1e-2) if step == 0 else (1e-1, 1e-1)
        assert rel(m["train/iit_loss"], GOLD[f"{mode}/train/iit_loss"][step]) < tol_iit, (step, m)
        assert rel(m["train/behavior_loss"], GOLD[f"{mode}/train/behavior_loss"][step]) < tol_beh, (step, m)
    ll.eval()
    with torch.no_grad():
        base, src = dev(R.synthetic_pvr(B, 300)), dev(R.synthetic_pvr(B, 301))
        state = pair.rng.bit_generator.state
        nodes.append(str(pair.sample_hl_name()))
        pair.rng.bit_generator.state = state
        ev = pair.run_eval_step(base, src, pair.loss_fn)
    assert nodes == GOLD[f"{mode}/nodes"].tolist()          # node selection: bit-exact
    # eval metrics after free-running training: counts may differ by a sample or two (see above); exact IIA
    # equality is asserted from synchronised weights in test_eval_step_counts_bit_exact_from_synced_weights
    assert abs(ev["val/IIA"] - float(GOLD[f"{mode}/eval/val/IIA"])) <= 0.1
    assert abs(ev["val/accuracy"] - float(GOLD[f"{mode}/eval/val/accuracy"])) <= 0.1
    assert rel(ev["val/iit_loss"], float(GOLD[f"{mode}/eval/val/iit_loss"])) < 1e-1
    sd = ll.state_dict()
    close_after_updates(sd["mod.fc.mod.weight"].cpu().numpy(), GOLD[f"{mode}/final/fc_weight"], "fc.weight")
    assert int(sd["mod.bn1.mod.num_batches_tracked"]) == int(GOLD[f"{mode}/final/num_batches_tracked"])


def test_get_corr_probe_moves_running_var_like_reference():
    """pvr_hl.get_corr runs a zeros batch in train mode: running_var -> 0.9 everywhere (reference behaviour)."""
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    ll, _, _ = get_alignment("mnist_pvr", {"input_shape": (1, 3, 56, 56)})
    sd = ll.state_dict()
    torch.testing.assert_close(sd["mod.layer2.mod.0.mod.bn1.mod.running_var"].cpu(), torch.full((128,), 0.9))
    assert int(sd["mod.bn1.mod.num_batches_tracked"]) == 1


@pytest.mark.parametrize("hp", [HOOK, "mod.layer1.mod.0.mod.conv1.hook_point"])
def test_sweep_matches_reference_golden(hp):
    """IIA sweep semantic (eval_causality.py:47-85 as intended) for 12 leaky nodes at two hook points."""
    from iit_b200 import ops
    pair = make_pair("q", task="pvr_leaky", hook=hp)
    pair.ll_model.eval()
    assert [k.name for k in pair.corr] == GOLD[f"sweep/{hp}/nodes"].tolist()
    base_cpu = R.synthetic_pvr(B, 400)
    base = dev(base_cpu)
    stats = []
    with torch.no_grad():
        for i, k in enumerate(pair.corr):
            src_cpu = R.synthetic_pvr_counterfactual(base_cpu, k.name, 500 + i)
            src = dev(src_cpu)
            hl, llo = pair.do_intervention(base, src, k.name)  # a str key works, as eval_causality.py:70 passes it
            if i == 0:
                assert np.array_equal(hl.cpu().numpy(), GOLD[f"sweep/{hp}/first/hl"])
                np.testing.assert_allclose(llo.cpu().numpy(), GOLD[f"sweep/{hp}/first/ll"], rtol=1e-3, atol=1e-4)
            counts, _ = ops.iia_count(llo, hl, changed=(src[1] != base[1]))
            hits, den = counts.tolist()
            stats.append(np.float32(hits) / (np.float32(den) + np.float32(1e-10)))
    np.testing.assert_allclose(np.array(stats, dtype=np.float64), GOLD[f"sweep/{hp}/stats"], rtol=0, atol=1e-7)


def _strict_pairs(mode="q"):
    from iit_b200.model_pairs import StrictIITModelPair
    pair = make_pair(mode, cls=StrictIITModelPair, targs={"strict_weight": 0.4})
    want_names = [n for n in R.ResNet18Oracle.conv_hook_points() if n != HOOK]
    assert [n.name for n in pair.nodes_not_in_circuit] == want_names
    oll = R.ResNet18Oracle(R.resnet18_state_dict(0))
    opair = R.OraclePair(oll, "pvr", R.pvr_corr(mode, HOOK, 256, 4), lr=1e-3, weights=(1.0, 1.0, 0.4),
                         strict_nodes=[(n, None) for n in want_names])
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    return pair, opair, oll


def _grad_rel_l2(ll, oll):
    num = den = 0.0
    for k, p in ll.named_parameters():
        g, og = p.grad.detach().cpu().double(), oll.p[k].grad.double()
        num += float(((g - og) ** 2).sum())
        den += float((og ** 2).sum())
    return (num / den) ** 0.5


@pytest.mark.parametrize("mode", ["q", "c"])
def test_teacher_forced_terms_match_oracle(mode):
    """Every loss term of the strict train step (IIT, strict, behaviour; strict_iit_model_pair.py:36-91),
    evaluated by both sides FROM THE SAME WEIGHTS (re-synchronised from the oracle before each term, the
    oracle then applies its own clip+Adam update): loss within 1e-3 relative (north_star bar), gradient
    within 1e-2 in relative L2 over all 11.2 M parameters. The strict term for the ResNet has no reference
    implementation; its oracle is the restatement with the same node definition."""
    pair, opair, oll = _strict_pairs(mode)
    ll = pair.ll_model
    opt = pair.make_optimizer()
    ll.train()
    worst = 0.0
    for step in range(2):
        b, s = R.synthetic_pvr(B, 100 + step), R.synthetic_pvr(B, 200 + step)
        bd, sd_ = dev(b), dev(s)
        hl_node = pair.sample_hl_name()
        assert str(hl_node) == opair.sample_hl_name()
        ll_node = pair.sample_ll_node()
        o_ll_node = opair.sample_ll_node()
        assert ll_node.name == o_ll_node[0]
        terms = [
            ("iit", lambda: pair.get_IIT_loss_over_batch(bd, sd_, hl_node, pair.loss_fn),
             lambda: opair.iit_loss(b, s, str(hl_node))),
            ("strict", lambda: pair.get_strict_loss_over_batch(bd, sd_, ll_node, pair.loss_fn) * 0.4,
             lambda: opair.loss_fn(R.ll_intervention(oll, b[0], s[0], [o_ll_node])[0], b[1]) * 0.4),
            ("behaviour", lambda: pair.get_behaviour_loss_over_batch(bd, pair.loss_fn),
             lambda: opair.loss_fn(oll(b[0]), b[1])),
        ]
        for name, mine_fn, oracle_fn in terms:
            ll.load_state_dict(oll.state_dict())       # same weights and BN buffers on both sides
            opt.zero_grad()
            lm = mine_fn()
            lm.backward()
            opair.opt.zero_grad()
            lo = oracle_fn()
            lo.backward()
            assert rel(lm.item(), lo.item()) < 1e-3, (step, name, lm.item(), lo.item())
            worst = max(worst, rel(lm.item(), lo.item()))
            e = _grad_rel_l2(ll, oll)
            assert e < 1e-2, (step, name, e)
            torch.nn.utils.clip_grad_norm_(oll.parameters(), 1.0)
            opair.opt.step()
    print(f"worst relative loss error from synchronised weights: {worst:.2e}")


def test_fused_update_matches_oracle_update():
    """One clip+Adam update through FlatAdam's fused kernels vs torch's clip_grad_norm_ + Adam on the oracle,
    from identical weights and (near-identical) gradients: parameter deltas agree in relative L2."""
    pair, opair, oll = _strict_pairs("q")
    ll = pair.ll_model
    opt = pair.make_optimizer()
    ll.train()
    b, s = R.synthetic_pvr(B, 100), R.synthetic_pvr(B, 200)
    before = {k: v.detach().clone() for k, v in oll.p.items()}
    node = pair.sample_hl_name()
    pair.step_on_loss(pair.get_IIT_loss_over_batch(dev(b), dev(s), node, pair.loss_fn), opt)
    opair.step_on_loss(opair.iit_loss(b, s, str(node)))
    num = den = 0.0
    for k, p in ll.named_parameters():
        dm = p.detach().cpu().double() - before[k].double()
        do = oll.p[k].detach().double() - before[k].double()
        num += float(((dm - do) ** 2).sum())
        den += float((do ** 2).sum())
    # |delta_i| ~ lr for every element on Adam's first step, so an element whose gradient sign differs (rounding-
    # level gradients) contributes (2 lr)^2: 0.15 relative L2 allows ~0.5 % of the 11.2 M elements to do so
    assert (num / den) ** 0.5 < 0.15, (num / den) ** 0.5


@pytest.mark.parametrize("mode", ["q", "c"])
def test_eval_step_counts_bit_exact_from_synced_weights(mode):
    """run_eval_step (iit_behavior_model_pair.py:102-141) after training: weights taken from the oracle's
    trajectory, then IIA and accuracy must be EQUAL (integer counts) and the loss within 1e-3."""
    from iit_b200.model_pairs import IITBehaviorModelPair
    pair = make_pair(mode, cls=IITBehaviorModelPair)
    oll = R.ResNet18Oracle(R.resnet18_state_dict(0))
    opair = R.OraclePair(oll, "pvr", R.pvr_corr(mode, HOOK, 256, 4), lr=1e-3)
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    for step in range(2):
        opair.train_step(R.synthetic_pvr(B, 100 + step), R.synthetic_pvr(B, 200 + step))
        pair.sample_hl_name()
    pair.ll_model.load_state_dict(oll.state_dict())
    pair.ll_model.eval()
    for i in range(3):
        b, s = R.synthetic_pvr(B, 300 + i), R.synthetic_pvr(B, 400 + i)
        want = opair.eval_step(b, s)
        with torch.no_grad():
            got = pair.run_eval_step(dev(b), dev(s), pair.loss_fn)
        assert got["val/IIA"] == want["val/IIA"] and got["val/accuracy"] == want["val/accuracy"], (got, want)
        assert rel(got["val/iit_loss"], want["val/iit_loss"]) < 1e-3


def test_all_hook_points_cache_matches_oracle():
    """run_with_cache without a filter (reference behaviour: every hook point) in eval mode."""
    from iit_b200.nn.resnet import ResNet18LL
    ll = ResNet18LL().to(DEV)
    ll.load_state_dict(R.resnet18_state_dict(0))
    ll.eval()
    oll = R.ResNet18Oracle(R.resnet18_state_dict(0)).eval()
    x = R.synthetic_pvr(4, 7)[0]
    ocache = {}
    with torch.no_grad():
        want = oll(x, cache=ocache)
        got, cache = ll.run_with_cache(x.to(DEV))
    assert set(cache) == set(ocache)
    np.testing.assert_allclose(got.cpu().numpy(), want.numpy(), rtol=1e-3, atol=1e-4)
    for name in ("mod.conv1.hook_point", "mod.maxpool.hook_point", HOOK, "mod.layer2.mod.0.mod.downsample.hook_point",
                 "mod.layer4.hook_point", "mod.avgpool.hook_point"):
        np.testing.assert_allclose(cache[name].cpu().numpy(), ocache[name].numpy(), rtol=1e-3, atol=1e-4, err_msg=name)


def test_bf16_path_close_to_fp32_oracle():
    """bf16 activations / fp32 accumulation. Stated separately from the fp32 bar: losses within 5e-2
    relative of the fp32 oracle on these 8-sample batches (bf16 has 8 mantissa bits; 20 conv+BN layers)."""
    pair = make_pair("q", dtype=torch.bfloat16)
    pair.ll_model.train()
    oll = R.ResNet18Oracle(R.resnet18_state_dict(0))
    opair = R.OraclePair(oll, "pvr", R.pvr_corr("q", HOOK, 256, 4), lr=1e-3)
    opt = pair.make_optimizer()
    b, s = R.synthetic_pvr(B, 100), R.synthetic_pvr(B, 200)
    want = opair.train_step(b, s)
    got = pair.run_train_step(dev(b), dev(s), pair.loss_fn, opt)
    for key in ("train/iit_loss", "train/behavior_loss"):
        assert rel(got[key], want[key]) < 5e-2, (key, got, want)


def test_cuda_graph_step_equals_eager_step():
    """training_args["cuda_graphs"]: each (term, node) variant replayed from a captured graph. Same kernels, same
    order, same inputs => the losses of a free-running 3-step trajectory must agree with the eager path to fp32
    rounding of the atomically accumulated BN sums re-rounded to bf16 (1e-3 at this 8-pair batch), and capture must leave model/optimizer state untouched."""
    from iit_b200.model_pairs import StrictIITModelPair
    res = {}
    for mode in ("eager", "graphs"):
        pair = make_pair("q", cls=StrictIITModelPair, dtype=torch.bfloat16,
                         targs={"strict_weight": 0.4, "cuda_graphs": mode == "graphs"})
        opt = pair.make_optimizer()
        pair.ll_model.train()
        b, s = dev(R.synthetic_pvr(B, 100)), dev(R.synthetic_pvr(B, 200))
        if mode == "graphs":
            before = {k: v.clone() for k, v in pair.ll_model.state_dict().items()}
            pair.prepare_graphs(b, s, pair.loss_fn, opt)
            after = pair.ll_model.state_dict()
            assert all(torch.equal(before[k], after[k]) for k in before), "prepare_graphs must not change the model"
            assert opt.step_count == 0
            assert len(pair._graphs) == 4 + 16 + 1
        out = []
        for step in range(3):
            b, s = dev(R.synthetic_pvr(B, 100 + step)), dev(R.synthetic_pvr(B, 200 + step))
            out.append(pair.run_train_step(b, s, pair.loss_fn, opt))
        res[mode] = out
        assert opt.step_count == 9
    for i, (e, g) in enumerate(zip(res["eager"], res["graphs"])):
        for k in e:  # bf16 free-running trajectories with atomically ordered BN sums (see the chaos note above)
            assert rel(g[k], e[k]) < (5e-2 if i == 0 else 2e-1), (i, k, e, g)
    assert rel(res["graphs"][0]["train/iit_loss"], res["eager"][0]["train/iit_loss"]) < 1e-3


@pytest.mark.parametrize("dtype,bar", [(torch.float32, 1e-3), (torch.bfloat16, 5e-2)])
def test_84px_default_image_size_matches_oracle(dtype, bar):
    """The reference's DEFAULT image size (pad 7: 84x84, task_loader.py:10,43; the contract size is 56): odd spatial sizes
    21 / 11 / 6 / 3 downstream, the hooked layer is [B,256,6,6] with 3x3 quadrant boxes. IIT and behaviour terms from
    synchronised weights against the oracle: losses within the bar (1e-3 fp32, 5e-2 bf16), fp32 gradients within 1e-2 in
    relative L2 over all parameters."""
    from iit_b200.model_pairs import IITBehaviorModelPair
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    ll, hl, corr = get_alignment("mnist_pvr", {"mode": "q", "hook_point": HOOK, "input_shape": (1, 3, 84, 84)}, dtype=dtype)
    ll.load_state_dict(R.resnet18_state_dict(0))
    pair = IITBehaviorModelPair(hl, ll, corr, training_args={"lr": 1e-3})
    oll = R.ResNet18Oracle(R.resnet18_state_dict(0))
    opair = R.OraclePair(oll, "pvr", R.pvr_corr("q", HOOK, 256, 6), lr=1e-3)
    opt = pair.make_optimizer()
    ll.train()
    b, s = R.synthetic_pvr(B, 100, size=84), R.synthetic_pvr(B, 200, size=84)
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
        assert rel(lm.item(), lo.item()) < bar, (name, lm.item(), lo.item())
        if dtype == torch.float32:
            e = _grad_rel_l2(ll, oll)
            assert e < 1e-2, (name, e)
