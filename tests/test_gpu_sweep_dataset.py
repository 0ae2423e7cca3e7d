"""SURVEY.md 8(f) rows 1 and 3 and BASELINE.json configs[4], on the GPU through the C-ABI:
device-side pair batches (bit-exact against the reference-order IITDataset), the counterfactual generator (quadrant
swap with the "different class" constraint) and the causality sweep against the CPU oracle on the SAME counterfactuals."""
import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import restate as R  # noqa: E402

DEV = "cuda"
HOOK = "mod.layer3.mod.1.mod.conv2.hook_point"


@pytest.mark.parametrize("dtype,shape", [(torch.float32, (37, 3, 8, 8)), (torch.int64, (50,)), (torch.int64, (23, 4)),
                                         (torch.uint8, (19, 7)), (torch.bfloat16, (16, 33))])
def test_gather_rows_is_index_select(dtype, shape):
    from iit_b200 import ops
    g = torch.Generator().manual_seed(0)
    src = (torch.randn(shape, generator=g) * 50).to(dtype)
    idx = torch.randint(0, shape[0], (41,), generator=g)
    out = ops.gather_rows(src.to(DEV), idx.to(DEV))
    assert torch.equal(out.cpu(), src[idx])
    assert ops.gather_rows(src.to(DEV), idx[:0].to(DEV)).shape[0] == 0
    ops.raise_on_index_error(DEV)
    ops.gather_rows(src.to(DEV), torch.tensor([shape[0]], device=DEV))
    with pytest.raises(IndexError):
        ops.raise_on_index_error(DEV)


def test_device_dataset_batches_equal_reference_order_batches():
    """Same seed, same torch RNG state -> the same (base, ablation) batches, bit for bit, as the per-sample IITDataset
    whose loader is the reference's (DataLoader(shuffle=True) + collate_fn stacking samples)."""
    from iit_b200.utils.iit_dataset import DeviceIITDataset, IITDataset
    x, y, v = R.synthetic_pvr(53, 3)
    samples = [(x[i], y[i], v[i]) for i in range(53)]
    ref = IITDataset(samples, samples, seed=5, device=DEV)
    dev = DeviceIITDataset((x.to(DEV), y.to(DEV), v.to(DEV)), (x.to(DEV), y.to(DEV), v.to(DEV)), seed=5, device=DEV)
    assert len(ref) == len(dev)
    assert [ref.pair_indices(i) for i in range(len(ref))] == [tuple(r) for r in dev.pair_table().tolist()]
    torch.manual_seed(11)
    want = list(ref.make_loader(16, 0))
    torch.manual_seed(11)
    got = list(dev.make_loader(16, 0))
    assert len(want) == len(got) == 4
    for (wb, wa), (gb, ga) in zip(want, got):
        for a, b in zip((*wb, *wa), (*gb, *ga)):
            assert a.shape == b.shape and a.dtype == b.dtype and torch.equal(a.cpu(), b.cpu())
    # every_combination: item -> (i // n, i % n)
    ec = DeviceIITDataset((x[:5].to(DEV), y[:5].to(DEV), v[:5].to(DEV)), (x[:3].to(DEV), y[:3].to(DEV), v[:3].to(DEV)),
                          every_combination=True)
    assert len(ec) == 15 and ec.pair_table().tolist()[7] == [2, 1]


def _bank(seed=0, per_class=6, q=28):
    from iit_b200.tasks.mnist_pvr.counterfactual import DigitBank
    g = torch.Generator().manual_seed(seed)
    labels = torch.arange(10).repeat(per_class)[torch.randperm(10 * per_class, generator=g)]
    images = torch.rand(10 * per_class, 1, q, q, generator=g)
    return DigitBank(images.to(DEV), labels.to(DEV), seed=seed), images, labels


@pytest.mark.parametrize("node,quad", [("hook_tl", (0, 0)), ("hook_tr", (0, 1)), ("hook_bl", (1, 0)), ("hook_br", (1, 1))])
def test_counterfactual_generator(node, quad):
    """dataset.py:125-176: exactly the node's quadrant is replaced, by a bank digit of a DIFFERENT class; the
    intermediate variable and the label follow (label = oracle HL forward of the new variables)."""
    from iit_b200.tasks.mnist_pvr.pvr_hl import hl_nodes
    bank, images, labels = _bank()
    base = R.synthetic_pvr(64, 7)
    x, y, v = (t.to(DEV) for t in base)
    nx, ny, nv = bank.patch_batch_at_hl(x, v, hl_nodes[node], y)
    i = quad[0] * 2 + quad[1]
    r0, c0 = quad[0] * 28, quad[1] * 28
    mask = torch.ones_like(x, dtype=torch.bool)
    mask[:, :, r0:r0 + 28, c0:c0 + 28] = False
    assert torch.equal(nx[mask], x[mask])
    assert bool((nv[:, i] != v[:, i]).all())
    others = [j for j in range(4) if j != i]
    assert torch.equal(nv[:, others], v[:, others])
    assert torch.equal(ny.cpu(), R.pvr_hl_forward(nv.cpu()))
    # every pasted quadrant is a bank digit whose label is the new variable
    sorted_imgs = bank.images.cpu()
    sorted_labs = bank.labels.cpu()
    for b in range(0, 64, 7):
        quad_px = nx[b, :, r0:r0 + 28, c0:c0 + 28].cpu()
        hits = [(sorted_imgs[m] == quad_px).all().item() for m in range(sorted_imgs.shape[0])]
        assert sum(hits) == 1 and int(sorted_labs[hits.index(True)]) == int(nv[b, i])
    # the draws are spread over the other classes and differ between calls
    nx2, _, nv2 = bank.patch_batch_at_hl(x, v, hl_nodes[node], y)
    assert not torch.equal(nv2, nv) and len(set(nv[:, i].tolist())) > 5
    from iit_b200 import ops
    ops.raise_on_index_error(DEV)


@pytest.mark.parametrize("task,n_nodes", [("mnist_pvr", 4), ("pvr_leaky", 12)])
def test_causality_sweep_matches_oracle_on_the_same_counterfactuals(task, n_nodes):
    """eval_causality.py:14-101 at two hook points, two batches, for the 4-node correspondence and for the 12-node LEAKY one
    the reference's script sweeps (eval_causality.py:102,138; pvr_check_leaky_hl.py:72-109: HL variable i interchanged while
    LL quadrant j is patched): the library's sweep (kernels, truncated source run, integer counts) against the oracle's
    restatement fed the very counterfactual tensors the sweep drew."""
    from iit_b200 import eval_causality
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    from iit_b200.tasks.mnist_pvr.counterfactual import DigitBank
    ll, _, _ = get_alignment("mnist_pvr", {"input_shape": (1, 3, 56, 56)}, device=DEV)
    sd = R.resnet18_state_dict(0)
    ll.load_state_dict(sd)
    oll = R.ResNet18Oracle(sd)
    oll.eval()
    bank, _, _ = _bank(seed=3)
    batches = [tuple(t.to(DEV) for t in R.synthetic_pvr(24, 40 + i)) for i in range(2)]
    hooks = ["mod.layer1.mod.0.mod.conv1.hook_point", HOOK]
    drawn = []
    inner = bank.patch_batch_at_hl

    def recording(batch, intermediate_vars, hl_node, labels=None):
        out = inner(batch, intermediate_vars, hl_node, labels)
        drawn.append(tuple(t.cpu() for t in out))
        return out

    bank.patch_batch_at_hl = recording
    stats = eval_causality.evaluate_model_on_ablations(ll, task, batches, {"batch_size": 24, "count_pairs": True}, bank,
                                                       hook_points=hooks)
    assert stats.pop("_pairs") == 2 * n_nodes * 2 * 24
    k = 0
    for hp in hooks:
        dim = {hooks[0]: 14, HOOK: 4}[hp]
        ch = {hooks[0]: 64, HOOK: 256}[hp]
        corr = R.pvr_corr("q", hp, ch, dim) if task == "mnist_pvr" else R.pvr_leaky_corr(hp, dim)
        assert len(corr) == n_nodes
        assert [str(n) for n in stats[hp].keys()] == list(corr.keys())
        for node_name, ll_nodes in corr.items():
            ratios = []
            for base in batches:
                src = drawn[k]
                k += 1
                b = tuple(t.cpu() for t in base)
                hl_out = R.pvr_hl_intervened(b[2], src[2], node_name)
                with torch.no_grad():
                    ll_out, _ = R.ll_intervention(oll, b[0], src[0], ll_nodes)
                ratios.append(R.causality_stat(ll_out, hl_out, b[1], src[1]))
            want = float(torch.stack(ratios).sum() / len(ratios))
            got = [v for n, v in stats[hp].items() if str(n) == node_name][0]
            assert abs(got - want) <= 1e-6, (hp, node_name, got, want)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_sweep_with_shared_base_prefix_equals_full_base_runs(dtype):
    """The sweep computes the base run's prefix (everything before the residual block that holds the hook point) once per
    batch and resumes each node's base run there: same statistics, bit for bit, as full base runs (eval mode)."""
    from iit_b200 import eval_causality
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    ll, _, _ = get_alignment("mnist_pvr", {"input_shape": (1, 3, 56, 56)}, device=DEV, dtype=dtype)
    ll.load_state_dict(R.resnet18_state_dict(0))
    batches = [tuple(t.to(DEV) for t in R.synthetic_pvr(32, 60 + i)) for i in range(2)]
    hooks = ["mod.conv1.hook_point", "mod.layer1.mod.0.mod.conv1.hook_point", "mod.layer2.mod.0.mod.conv2.hook_point",
             "mod.layer3.mod.1.mod.conv2.hook_point", "mod.layer4.mod.1.mod.conv1.hook_point"]
    assert ll.resume_point(hooks[0]) is None
    assert ll.resume_point(hooks[3]) == (5, "mod.layer3.mod.0.hook_point")
    assert ll.resume_point("mod.layer2.mod.0.mod.conv1.hook_point") == (2, "mod.layer1.hook_point")
    res = []
    for share in (True, False):
        bank, _, _ = _bank(seed=9)
        res.append(eval_causality.evaluate_model_on_ablations(ll, "pvr_leaky", batches, {"batch_size": 32}, bank, hook_points=hooks,
                                                              share_base_prefix=share))
    for hp in hooks:
        assert {str(k): v for k, v in res[0][hp].items()} == {str(k): v for k, v in res[1][hp].items()}, hp
