"""Interventions at hook points other than convolution outputs, and the small contracts around the hook machinery.

The reference clones the hooked tensor before it patches (iit/model_pairs/base_model_pair.py:157-161), so it trains with a
correspondence on ANY hook point. Here the patch is in place only where the model vouches that no operator kept the tensor
for its backward (HookPoint.inplace_ok: convolution outputs, the transformer's hook points); BatchNorm / ReLU / block outputs
are cloned first. These tests train through such hook points against the CPU oracle."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import restate as R  # noqa: E402

DEV = "cuda"
B = 8


def dev(inp):
    return tuple(t.to(DEV) for t in inp)


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-12)


def _grad_rel_l2(ll, oll):
    num = den = 0.0
    for k, p in ll.named_parameters():
        g, og = p.grad.detach().cpu().double(), oll.p[k].grad.double()
        num += float(((g - og) ** 2).sum())
        den += float((og ** 2).sum())
    return (num / den) ** 0.5


@pytest.mark.parametrize("hp", ["mod.layer3.mod.1.hook_point",                 # block output (after the residual ReLU)
                                "mod.layer3.mod.1.mod.relu.hook_point",       # the block's shared ReLU: fires twice
                                "mod.layer2.mod.0.mod.bn1.hook_point",        # a BatchNorm output (unfused path)
                                "mod.maxpool.hook_point"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_train_step_with_correspondence_on_saved_output_hook_points(hp, dtype):
    """IIT term + strict term (whole-tensor patch at `hp`) + behaviour, each from synchronised weights: the backward must
    run (the in-place patch used to invalidate the tensor its producer saved) and agree with the oracle."""
    from iit_b200.model_pairs import StrictIITModelPair
    from iit_b200.model_pairs.nodes import LLNode
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    from iit_b200.utils.index import Ix
    torch.set_num_threads(max(1, os.cpu_count() or 1))
    ll, hl, corr = get_alignment("mnist_pvr", {"mode": "q", "hook_point": hp, "input_shape": (1, 3, 56, 56)}, dtype=dtype)
    ll.load_state_dict(R.resnet18_state_dict(0))
    pair = StrictIITModelPair(hl, ll, corr, training_args={"lr": 1e-3, "nodes_not_in_circuit": [LLNode(hp, Ix[[None]])]})
    opt = pair.make_optimizer()
    ll.train()
    oll = R.ResNet18Oracle(R.resnet18_state_dict(0))
    with torch.no_grad():
        ocache = {}
        oll(torch.zeros(1, 3, 56, 56), cache=ocache, cache_filter={hp})   # the corr probe of get_alignment (moves BN stats)
    _, ch, dim, _ = ocache[hp].shape
    opair = R.OraclePair(oll, "pvr", R.pvr_corr("q", hp, ch, dim), lr=1e-3, strict_nodes=[(hp, None)])
    b, s = R.synthetic_pvr(B, 100), R.synthetic_pvr(B, 200)
    bd, sd_ = dev(b), dev(s)
    hl_node = pair.sample_hl_name()
    assert str(hl_node) == opair.sample_hl_name()
    ll_node = pair.sample_ll_node()
    # bf16 at this 8-pair batch, stated separately: loss 2e-2 (measured 2e-3 .. 3.3e-3); the gradient of the randomly
    # initialised network is sensitive to activation rounding itself (tests/test_bf16_rounding_model.py), measured 0.40
    loss_bar, grad_bar = (1e-3, 1e-2) if dtype == torch.float32 else (2e-2, 6e-1)
    for name, mine_fn, oracle_fn in [
            ("iit", lambda: pair.get_IIT_loss_over_batch(bd, sd_, hl_node, pair.loss_fn), lambda: opair.iit_loss(b, s, str(hl_node))),
            ("strict", lambda: pair.get_strict_loss_over_batch(bd, sd_, ll_node, pair.loss_fn),
             lambda: opair.loss_fn(R.ll_intervention(oll, b[0], s[0], [(hp, None)])[0], b[1])),
            ("behaviour", lambda: pair.get_behaviour_loss_over_batch(bd, pair.loss_fn), lambda: opair.loss_fn(oll(b[0]), b[1]))]:
        ll.load_state_dict(oll.state_dict())
        opt.zero_grad()
        lm = mine_fn()
        lm.backward()
        opair.opt.zero_grad()
        lo = oracle_fn()
        lo.backward()
        e = _grad_rel_l2(ll, oll)
        print(f"{hp} {dtype} {name}: loss rel {rel(lm.item(), lo.item()):.2e}, grad rel-L2 {e:.2e}")
        assert rel(lm.item(), lo.item()) < loss_bar, (hp, name, lm.item(), lo.item())
        assert e < grad_bar, (hp, name, e)
    # and the whole step runs with the fused optimizer
    m = pair.run_train_step(bd, sd_, pair.loss_fn, opt)
    assert all(np.isfinite(v) for v in m.values()), m


def test_stop_when_cached_keeps_last_firing_of_a_shared_relu():
    """A BasicBlock's ReLU hook point fires after bn1 and again after the residual add; a cache holds the LAST value
    (transformer_lens and the reference wrapper). The early-stopping source run must not stop at the first firing."""
    from iit_b200.nn.resnet import ResNet18LL
    ll = ResNet18LL().to(DEV)
    ll.load_state_dict(R.resnet18_state_dict(0))
    ll.eval()
    x = R.synthetic_pvr(4, 3)[0].to(DEV)
    name = "mod.layer1.mod.0.mod.relu.hook_point"
    with torch.no_grad():
        _, full = ll.run_with_cache(x, names_filter=[name])
        out, stop = ll.run_with_cache(x, names_filter=[name], stop_when_cached=True)
        _, block = ll.run_with_cache(x, names_filter=["mod.layer1.mod.0.hook_point"])
    assert torch.equal(full[name], stop[name])
    assert torch.equal(full[name], block["mod.layer1.mod.0.hook_point"])   # = the block output
    oc = {}
    with torch.no_grad():
        R.ResNet18Oracle(R.resnet18_state_dict(0)).eval()(x.cpu(), cache=oc, cache_filter={name})
    np.testing.assert_allclose(stop[name].cpu().numpy(), oc[name].numpy(), rtol=1e-3, atol=1e-4)
    # a hook point that fires once still stops early (the forward is abandoned: no output)
    with torch.no_grad():
        out2, c2 = ll.run_with_cache(x, names_filter=["mod.layer1.mod.0.mod.conv1.hook_point"], stop_when_cached=True)
    assert out2 is None and "mod.layer1.mod.0.mod.conv1.hook_point" in c2


def test_hl_pvr_out_of_range_pointer_raises_like_indexing():
    from iit_b200 import ops
    v = torch.tensor([[3, 1, 2, 4], [-10, 5, 6, 7], [9, 1, 2, 3]], device=DEV)
    assert ops.hl_pvr(v, None, -1).tolist() == [1, 5, 3]
    ops.raise_on_index_error(DEV)
    bad = torch.tensor([[3, 1, 2, 4], [10, 5, 6, 7], [-11, 1, 2, 3]], device=DEV)
    assert ops.hl_pvr(bad, None, -1).tolist() == [1, -1, -1]
    with pytest.raises(IndexError):
        ops.raise_on_index_error(DEV)
    ops.raise_on_index_error(DEV)   # the flag was cleared


def test_hl_cache_is_available_after_do_intervention():
    """reference base_model_pair.py:80: `self.hl_cache` holds the HL source run after an intervention."""
    from iit_b200.model_pairs import IITBehaviorModelPair
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    ll, hl, corr = get_alignment("mnist_pvr", {"input_shape": (1, 3, 56, 56)})
    pair = IITBehaviorModelPair(hl, ll, corr, training_args={"lr": 1e-3})
    b, s = dev(R.synthetic_pvr(B, 100)), dev(R.synthetic_pvr(B, 200))
    ll.eval()
    with torch.no_grad():
        pair.do_intervention(b, s, "hook_bl")
    assert set(pair.hl_cache.keys()) >= {"hook_tl", "hook_tr", "hook_bl", "hook_br"}
    assert torch.equal(pair.hl_cache["hook_bl"], s[2][:, 2])


def test_check_causal_effect_all_nodes_runs_on_the_ioi_model():
    """node_type='a' builds nodes on the default `attn.hook_result` suffix (iit/utils/node_picker.py:11-14). With
    use_attn_result=False that hook point never fires, in transformer_lens and here: the head nodes report no effect and
    the MLP nodes (mlp.hook_post) are ablated."""
    from iit_b200.model_pairs import IOI_ModelPair
    from iit_b200.tasks.ioi import IOI_HL, corr, make_ll_model
    from iit_b200.utils.eval_ablations import check_causal_effect
    from iit_b200.utils.iit_dataset import IITDataset
    V = 101
    sd, data, names = R.ioi_ablation_fixture(0.0)
    ll = make_ll_model(device=DEV, d_mlp=128, d_vocab=V, n_ctx=32)
    ll.load_state_dict(sd)
    ll.eval()
    pair = IOI_ModelPair(IOI_HL(V, names).to(DEV), ll, corr, training_args={"next_token": False})
    ds = IITDataset(data, data, seed=0, device=DEV)
    torch.manual_seed(0)
    res = check_causal_effect(pair, ds, batch_size=16, node_type="a")
    heads = [k for k in res if k.name.endswith("attn.hook_result")]
    mlps = [k for k in res if k.name.endswith("mlp.hook_post")]
    assert len(heads) == 24 and len(mlps) == 6
    # a hook that never fires leaves the run unablated: every head node reports the same (no-ablation) value
    assert len({float(res[k]) for k in heads}) == 1
    assert all(float(res[k]) != float(res[heads[0]]) for k in mlps)
