"""Pins oracle/restate.py (the CPU checker) against
  (1) the reference's own literal test vectors (iit/tasks/ioi/test_ioi.py:10-45, tests/test_index.py,
      tests/test_node_picker.py:6-29), and
  (2) tests/golden/*.npz -- outputs of the UNMODIFIED reference run by oracle/make_golden.py."""
import os

import numpy as np
import pytest
import torch

from oracle import restate as R

GOLD = os.path.join(os.path.dirname(__file__), "golden")


@pytest.fixture(scope="module")
def pvr():
    return np.load(os.path.join(GOLD, "pvr.npz"))


@pytest.fixture(scope="module")
def ioi():
    return np.load(os.path.join(GOLD, "ioi.npz"))


def nonzero_values(a):
    return torch.cat((a.nonzero(), a[a != 0][:, None]), dim=-1)


# ---- (1) reference literal vectors --------------------------------------------------------
def test_ref_duplicate_head():
    a = R.ioi_duplicate_head(torch.tensor([[3, 1, 4, 1, 5, 9, 2, 6, 5]]))
    assert a.equal(torch.tensor([[-1, -1, -1, 1, -1, -1, -1, -1, 4]]))


def test_ref_s_inhibition_head():
    a = R.ioi_s_inhibition_head(torch.tensor([[3, 1, 4, 1, 5, 9, 2, 6, 5]]),
                                torch.tensor([[-1, -1, -1, 1, -1, -1, -1, -1, 4]]))
    assert a.equal(torch.tensor([[-1, -1, -1, 1, -1, -1, -1, -1, 5]]))


def test_ref_name_mover_head():
    a = R.ioi_name_mover_head(torch.tensor([[1, 2, 10, 20]]), torch.tensor([[-1, 20, 10, -1]]),
                              torch.tensor([10, 20, 30]), 21)
    assert nonzero_values(a[0]).equal(torch.tensor(
        [[1., 20., -15.], [2., 10., -5.], [2., 20., -15.], [3., 10., -5.], [3., 20., -5.]]))


def test_ref_ioi_hl():
    a = R.ioi_hl_forward(torch.tensor([[3, 10, 4, 10, 5, 9, 2, 6, 5]]), torch.tensor([10, 20, 30]), 21)
    assert nonzero_values(a[0]).equal(torch.tensor(
        [[1., 10., 10.], [2., 10., 10.], [3., 10., 5.], [4., 10., 5.], [5., 10., 5.], [6., 10., 5.],
         [7., 10., 5.], [8., 5., -15.], [8., 10., 5.]]))


def test_ref_index_intersects():
    s = slice(None)
    assert R.specs_intersect((s, s, s, s), (s, s, 0, s))
    assert not R.specs_intersect((s, s, 0, s), (s, s, slice(2, 3), s))
    assert not R.specs_intersect((s, s, 2, s), (s, s, 1, s))
    with pytest.raises(ValueError):
        R.specs_intersect((s, s, 1), (s, s, s, s))


def test_ref_get_all_nodes():
    s = slice(None)
    got = R.all_nodes(2, 4, {"attn": "attn.hook_result", "mlp": "mlp.hook_post"})
    want = []
    for l in range(2):
        want += [(f"blocks.{l}.attn.hook_result", (s, s, h, s)) for h in range(4)]
        want.append((f"blocks.{l}.mlp.hook_post", None))
    assert got == want


# ---- (2) golden fixtures from the reference run -------------------------------------------
HOOK = "mod.layer3.mod.1.mod.conv2.hook_point"
B = int(np.load(os.path.join(GOLD, "pvr.npz"))["B"])


def test_pvr_hook_names(pvr):
    ll = R.ResNet18Oracle(R.resnet18_state_dict(0))
    cache = {}
    ll.eval()
    ll(torch.zeros(1, 3, 56, 56), cache=cache)
    assert set(cache.keys()) == set(pvr["hook_points"].tolist())
    assert len(cache) == 67
    assert R.ResNet18Oracle.conv_hook_points() == pvr["conv_hook_points"].tolist()
    assert set(R.resnet18_state_dict(0).keys()) == set(pvr["state_dict_keys"].tolist())


@pytest.mark.parametrize("mode", ["q", "c"])
def test_pvr_do_intervention_and_train(pvr, mode):
    torch.set_num_threads(4)
    ll = R.ResNet18Oracle(R.resnet18_state_dict(0))
    corr = R.pvr_corr(mode, HOOK, 256, 4)
    pair = R.OraclePair(ll, "pvr", corr, lr=1e-3)
    base, src = R.synthetic_pvr(B, 100), R.synthetic_pvr(B, 200)
    ll.train()
    sd0 = ll.state_dict()
    for k in corr:
        hl, llo = pair.do_intervention(base, src, k)
        assert np.array_equal(hl.numpy(), pvr[f"{mode}/di/{k}/hl"])  # integer: bit-exact
        np.testing.assert_allclose(llo.detach().numpy(), pvr[f"{mode}/di/{k}/ll"], rtol=1e-4, atol=1e-5)
    ll.b.update({k: v.clone() for k, v in sd0.items() if k in ll.b})
    nodes = []
    for step in range(3):
        base, src = R.synthetic_pvr(B, 100 + step), R.synthetic_pvr(B, 200 + step)
        m = pair.train_step(base, src)
        nodes.append(m["hl_node"])
        # Tight only before the first update: a free-running fp32 trajectory through batch-stat BN + ReLU +
        # Adam is chaotic at rounding level -- this very test, run on the GPU box's host CPU, leaves the
        # fixture (generated in the build container) by > 1e-4 after one update and by percents at step 2.
        t_iit, t_beh = (1e-4, 1e-2) if step == 0 else (1e-1, 1e-1)
        np.testing.assert_allclose(m["train/iit_loss"], pvr[f"{mode}/train/iit_loss"][step], rtol=t_iit)
        np.testing.assert_allclose(m["train/behavior_loss"], pvr[f"{mode}/train/behavior_loss"][step], rtol=t_beh)
    base, src = R.synthetic_pvr(B, 300), R.synthetic_pvr(B, 301)
    ev = pair.eval_step(base, src)
    nodes.append(ev["hl_node"])
    assert nodes == pvr[f"{mode}/nodes"].tolist()  # node selection: bit-exact
    assert abs(ev["val/IIA"] - float(pvr[f"{mode}/eval/val/IIA"])) <= 0.1
    assert abs(ev["val/accuracy"] - float(pvr[f"{mode}/eval/val/accuracy"])) <= 0.1
    np.testing.assert_allclose(ev["val/iit_loss"], float(pvr[f"{mode}/eval/val/iit_loss"]), rtol=1e-1)
    sd = ll.state_dict()
    d = np.abs(sd["mod.fc.mod.weight"].numpy() - pvr[f"{mode}/final/fc_weight"])
    assert d.mean() < 1e-3 and d.max() < 2e-2
    assert int(sd["mod.bn1.mod.num_batches_tracked"]) == int(pvr[f"{mode}/final/num_batches_tracked"])


def test_pvr_cache_is_source_activation(pvr):
    ll = R.ResNet18Oracle(R.resnet18_state_dict(0))
    ll.train()
    base, src = R.synthetic_pvr(B, 100), R.synthetic_pvr(B, 200)
    _, cache = R.ll_intervention(ll, base[0], src[0], R.pvr_corr("q", HOOK, 256, 4)["hook_tr"])
    np.testing.assert_allclose(cache[HOOK][:4].numpy(), pvr["q/di/hook_tr/cache_hook"], rtol=1e-4, atol=1e-5)


@pytest.mark.parametrize("hp", [HOOK, "mod.layer1.mod.0.mod.conv1.hook_point"])
def test_pvr_sweep(pvr, hp):
    ll = R.ResNet18Oracle(R.resnet18_state_dict(0)).eval()
    dim = 4 if hp == HOOK else 14
    corr = R.pvr_leaky_corr(hp, dim)
    assert list(corr.keys()) == pvr[f"sweep/{hp}/nodes"].tolist()
    base = R.synthetic_pvr(B, 400)
    stats = []
    with torch.no_grad():
        for i, k in enumerate(corr):
            src = R.synthetic_pvr_counterfactual(base, k, 500 + i)
            llo, _ = R.ll_intervention(ll, base[0], src[0], corr[k])
            hlo = R.pvr_hl_intervened(base[2], src[2], k)
            if i == 0:
                assert np.array_equal(hlo.numpy(), pvr[f"sweep/{hp}/first/hl"])
                np.testing.assert_allclose(llo.numpy(), pvr[f"sweep/{hp}/first/ll"], rtol=1e-4, atol=1e-5)
            stats.append(R.causality_stat(llo, hlo, base[1], src[1]).item())
    np.testing.assert_allclose(stats, pvr[f"sweep/{hp}/stats"], rtol=0, atol=1e-7)


def _ioi_pair():
    V = 101
    ll = R.TransformerOracle(R.transformer_state_dict(0, d_mlp=128, d_vocab=V, n_ctx=32), 6, 4, 16)
    _, names = R.synthetic_ioi(8, 100, d_vocab=V, n_names=20)
    corr = R.ioi_corr()
    strict = R.nodes_not_in_circuit(6, 4, corr, {"attn": "attn.hook_z", "mlp": "mlp.hook_post"})
    pair = R.OraclePair(ll, "ioi", corr, lr=1e-3, weights=(1.0, 1.0, 0.4), strict_nodes=strict,
                        names=names, d_vocab=V)
    return pair, strict, V


def _node_str(n):
    name, spec = n
    if spec is None:
        return f"{name}|[:]"
    return f"{name}|[" + ", ".join(":" if s == slice(None) else str(s) for s in spec) + "]"


def test_ioi_nodes(ioi):
    pair, strict, V = _ioi_pair()
    assert [_node_str(n) for n in strict] == ioi["nodes_not_in_circuit"].tolist()
    assert len(strict) == 8
    alln = R.all_nodes(6, 4, {"attn": "attn.hook_z", "mlp": "mlp.hook_post"})
    assert [_node_str(n) for n in alln] == ioi["all_nodes"].tolist()


def test_ioi_do_intervention_and_train(ioi):
    torch.set_num_threads(4)
    pair, strict, V = _ioi_pair()
    assert np.array_equal(pair.names.numpy(), ioi["names"])
    base, _ = R.synthetic_ioi(8, 100, d_vocab=V, n_names=20)
    src, _ = R.synthetic_ioi(8, 200, d_vocab=V, n_names=20)
    np.testing.assert_array_equal(R.ioi_hl_forward(base[0], pair.names, V).numpy(), ioi["hl/base_full"])
    for k in pair.corr:
        hl, llo = pair.do_intervention(base, src, k)
        np.testing.assert_array_equal(hl[:, -1, :].numpy(), ioi[f"di/{k}/hl_last"])
        np.testing.assert_allclose(llo[:, -1, :].detach().numpy(), ioi[f"di/{k}/ll_last"], rtol=1e-4, atol=1e-5)
        np.testing.assert_allclose(pair.iit_loss(base, src, k).item(), float(ioi[f"di/{k}/iit_loss"]), rtol=1e-4)
    seq = []
    for step in range(3):
        base, _ = R.synthetic_ioi(8, 100 + step, d_vocab=V, n_names=20)
        src, _ = R.synthetic_ioi(8, 200 + step, d_vocab=V, n_names=20)
        m = pair.train_step(base, src, strict=True)
        seq += [m["hl_node"], _node_str(m["ll_node"])]
        np.testing.assert_allclose(m["train/iit_loss"], ioi["train/iit_loss"][step], rtol=1e-4)
        np.testing.assert_allclose(m["train/strict_loss"], ioi["train/strict_loss"][step], rtol=1e-4)
        np.testing.assert_allclose(m["train/behavior_loss"], ioi["train/behavior_loss"][step], rtol=1e-4)
    assert seq == ioi["choice_sequence"].tolist()
    base, _ = R.synthetic_ioi(8, 300, d_vocab=V, n_names=20)
    src, _ = R.synthetic_ioi(8, 301, d_vocab=V, n_names=20)
    ev = pair.eval_step(base, src)
    assert ev["val/IIA"] == float(ioi["eval/val/IIA"])
    np.testing.assert_array_equal(ev["val/per_token_accuracy"], ioi["eval/val/per_token_accuracy"].astype(np.float32))
    np.testing.assert_allclose(ev["val/iit_loss"], float(ioi["eval/val/iit_loss"]), rtol=1e-3)
    np.testing.assert_allclose(pair.ll.p["blocks.3.mlp.W_in"][:4, :8].detach().numpy(), ioi["final/b3_W_in"],
                               rtol=1e-3, atol=1e-6)


# ------------------------------------------------------------------------------------------ ablation evaluations
def _parse_node(key):
    """'blocks.1.attn.hook_z|[:, :, 2, :]' -> (name, index spec) of the oracle."""
    name, idx = key.split("|")
    parts = [p.strip() for p in idx.strip("[]").split(",")]
    spec = tuple(None if p == ":" else int(p) for p in parts)
    return name, (None if spec == (None,) else spec)


def test_ablation_evaluations_match_reference_golden():
    """oracle/restate.py's restatement of iit/utils/eval_ablations.py + eval_metrics.py against the fixtures the
    unmodified reference produced (oracle/make_golden.py:golden_ioi_ablations): resample-ablation KL per node, the mean
    cache, mean- / zero-ablated logits, and the accuracy-based statistics."""
    g = np.load(os.path.join(os.path.dirname(__file__), "golden", "ioi_ablations.npz"))
    sd, data, names = R.ioi_ablation_fixture(0.0)
    ll = R.TransformerOracle(sd, 6, 4, 16).eval()
    for nt in ("n", "c"):
        nodes = [_parse_node(k) for k in g[f"resample_kl/{nt}/nodes"].tolist()]
        torch.manual_seed(0)
        got = R.resample_kl_effects(ll, names, 101, data, nodes, 16)
        np.testing.assert_allclose(got, g[f"resample_kl/{nt}/values"], rtol=1e-4)
    torch.manual_seed(0)
    mc = R.mean_cache(ll, data, ["blocks.1.attn.hook_z", "blocks.2.mlp.hook_post", "blocks.3.mlp.hook_post"], 16)
    for name in ("blocks.1.attn.hook_z", "blocks.2.mlp.hook_post"):
        np.testing.assert_allclose(mc[name].numpy(), g[f"mean_cache/{name}"], rtol=1e-4, atol=1e-5)
    x8 = torch.stack([d[0] for d in data[:8]])
    for key in ("blocks.1.attn.hook_z|[:, :, 2, :]", "blocks.3.mlp.hook_post|[:]"):
        node = _parse_node(key)
        for tag, mean in (("mean", mc), ("zero", None)):
            out = R.ablated_logits(ll, x8, node, mean)[:, -1, :]
            np.testing.assert_allclose(out.numpy(), g[f"ablated_logits/{key}/{tag}"], rtol=1e-3, atol=1e-4)
    sd4, data4, names4 = R.ioi_ablation_fixture(4.0)
    ll4 = R.TransformerOracle(sd4, 6, 4, 16).eval()
    for tag in ("mean", "zero"):
        for nt in ("n", "c"):
            nodes = [_parse_node(k) for k in g[f"ablate/{tag}/{nt}/nodes"].tolist()]
            torch.manual_seed(0)
            mean = R.mean_cache(ll4, data4, sorted({n for n, _ in nodes}), 16) if tag == "mean" else None
            got = R.ablation_effects(ll4, names4, 101, data4, nodes, 16, mean)
            np.testing.assert_allclose(got, g[f"ablate/{tag}/{nt}/values"], atol=1e-6)
