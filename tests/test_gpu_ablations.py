"""SURVEY.md 8(f) row 2 on the GPU: iit_b200.utils.eval_ablations / eval_metrics / eval_datasets against the fixtures
produced by the UNMODIFIED reference (tests/golden/ioi_ablations.npz, oracle/make_golden.py:golden_ioi_ablations) and
against the CPU oracle. Continuous quantities (KL, mean cache, ablated logits) to 1e-3; counts-based statistics exactly."""
import os

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from oracle import restate as R  # noqa: E402

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "ioi_ablations.npz"))
DEV = "cuda"
V = 101


def make_pair(bias):
    from iit_b200.model_pairs import IOI_ModelPair
    from iit_b200.tasks.ioi import IOI_HL, corr, make_ll_model
    sd, data, names = R.ioi_ablation_fixture(bias)
    ll = make_ll_model(device=DEV, d_mlp=128, d_vocab=V, n_ctx=32)
    ll.load_state_dict(sd)
    ll.eval()
    pair = IOI_ModelPair(IOI_HL(V, names).to(DEV), ll, corr, training_args={"next_token": False})
    return pair, data


def key(n):
    return f"{n.name}|{n.index}"


@pytest.mark.parametrize("rows,classes", [(7, 10), (5, 101), (3, 50257)])
def test_kl_div_kernel(rows, classes):
    """iit_kl_div_rows vs the reference formula (eval_metrics.py:4-20) for logits / probability inputs, zeros in b."""
    from iit_b200.utils.eval_metrics import kl_div
    from iit_b200.utils.index import Ix
    g = torch.Generator().manual_seed(rows)
    a, b = torch.randn(rows, classes, generator=g) * 3, torch.randn(rows, classes, generator=g) * 3
    pb = F.softmax(b, -1)
    pb[:, ::3] = 0
    pb = pb / pb.sum(-1, keepdim=True)
    for aa, bb in ((a, b), (F.softmax(a, -1), b), (a, pb)):
        want = R.kl_div_rows(aa, bb)
        got = kl_div(aa.to(DEV), bb.to(DEV), Ix[[None]])
        torch.testing.assert_close(got.cpu(), want, rtol=1e-4, atol=1e-5)
    got3 = kl_div(a.view(rows, 1, classes).expand(rows, 4, classes).to(DEV), b.view(rows, 1, classes).expand(rows, 4, classes).to(DEV),
                  Ix[:, -1])
    torch.testing.assert_close(got3.cpu(), R.kl_div_rows(a, b), rtol=1e-4, atol=1e-5)


def test_resample_ablation_kl_matches_reference_golden():
    from iit_b200.utils import eval_ablations as ea
    from iit_b200.utils.iit_dataset import IITDataset
    pair, data = make_pair(0.0)
    ds = IITDataset(data, data, seed=0, device=DEV)
    for nt in ("n", "c"):
        torch.manual_seed(0)
        got = {key(n): v for n, v in ea.check_causal_effect(pair, ds, batch_size=16, node_type=nt).items()}
        want = dict(zip(GOLD[f"resample_kl/{nt}/nodes"].tolist(), GOLD[f"resample_kl/{nt}/values"].tolist()))
        assert set(got) == set(want)
        for k in want:
            assert abs(got[k] - want[k]) <= 1e-3 * abs(want[k]), (nt, k, got[k], want[k])


def test_mean_cache_and_ablation_hooks_match_reference_golden():
    from iit_b200.model_pairs.nodes import LLNode
    from iit_b200.utils import eval_ablations as ea
    from iit_b200.utils.eval_datasets import IITUniqueDataset
    from iit_b200.utils.index import Ix
    pair, data = make_pair(0.0)
    uni = IITUniqueDataset(data, data, seed=0, device=DEV)
    torch.manual_seed(0)
    mc = ea.get_mean_cache(pair, uni, batch_size=16)
    for name in ("blocks.1.attn.hook_z", "blocks.2.mlp.hook_post", "blocks.0.hook_resid_pre"):
        np.testing.assert_allclose(mc[name].cpu().numpy(), GOLD[f"mean_cache/{name}"], rtol=1e-3, atol=1e-5)
    x8 = torch.stack([d[0] for d in data[:8]]).to(DEV)
    for node in (LLNode("blocks.1.attn.hook_z", Ix[:, :, 2, :]), LLNode("blocks.3.mlp.hook_post", Ix[[None]])):
        for tag, use_mean in (("mean", True), ("zero", False)):
            hook = ea.make_ablation_hook(node, mc, use_mean)
            with torch.no_grad():
                out = pair.ll_model.run_with_hooks(x8, fwd_hooks=[(node.name, hook)])[:, -1, :]
            np.testing.assert_allclose(out.cpu().numpy(), GOLD[f"ablated_logits/{key(node)}/{tag}"], rtol=1e-3, atol=1e-4)


def test_ablation_statistics_match_reference_golden_and_oracle():
    from iit_b200.utils import eval_ablations as ea
    from iit_b200.utils.eval_datasets import IITUniqueDataset
    pair, data = make_pair(4.0)
    uni = IITUniqueDataset(data, data, seed=0, device=DEV)
    for tag in ("mean", "zero"):
        torch.manual_seed(0)
        mc = ea.get_mean_cache(pair, uni, batch_size=16) if tag == "mean" else None
        for nt in ("n", "c"):
            got = {key(n): v for n, v in ea.check_causal_effect_on_ablation(pair, uni, batch_size=16, node_type=nt,
                                                                             mean_cache=mc).items()}
            want = dict(zip(GOLD[f"ablate/{tag}/{nt}/nodes"].tolist(), GOLD[f"ablate/{tag}/{nt}/values"].tolist()))
            assert set(got) == set(want)
            for k in want:
                assert abs(got[k] - want[k]) <= 1e-6, (tag, nt, k, got[k], want[k])
    # a case where the statistic is not zero: zero-ablating the whole MLP of every layer, against the oracle
    sd, _, names = R.ioi_ablation_fixture(4.0)
    oll = R.TransformerOracle(sd, 6, 4, 16).eval()
    nodes = [(f"blocks.{i}.mlp.hook_post", None) for i in range(6)] + [(f"blocks.{i}.attn.hook_z", None) for i in range(6)]
    torch.manual_seed(3)
    want = R.ablation_effects(oll, names, V, data, nodes, 16, None)
    from iit_b200.model_pairs.nodes import LLNode
    from iit_b200.utils.index import Ix
    results = {LLNode(n, Ix[[None]]): 0 for n, _ in nodes}
    torch.manual_seed(3)
    loader = uni.make_loader(16, 0)
    with torch.no_grad():
        for base_in in loader:
            for node in results:
                ea.ablate_node(pair, base_in, node, results, ea.make_ablation_hook(node, None, False))
    got = [v / len(loader) for v in results.values()]
    np.testing.assert_allclose(got, want, atol=1e-6)
