"""TEST INFRASTRUCTURE (oracle): generate tests/golden/*.npz by RUNNING THE UNMODIFIED REFERENCE.

Run in the build container only (needs /root/reference):  python oracle/make_golden.py
The reference's own classes (iit.model_pairs.*, iit.tasks.*, iit.utils.*) are imported from
/root/reference through oracle/ref_env.py (transformer_lens shim + synthetic MNIST files) and driven
on the seeded synthetic pairs of oracle/restate.py. Outputs are small fp32/int64 arrays that pin
  * oracle/restate.py (tests/test_oracle_golden.py, CPU) and
  * the CUDA path (tests/test_gpu_*.py, on the B200 box where the reference does not exist).
"""
from __future__ import annotations

import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
REPO = os.path.dirname(HERE)
sys.path.insert(0, REPO)

from oracle import ref_env, restate  # noqa: E402

OUT = os.path.join(REPO, "tests", "golden")


def to_np(d):
    out = {}
    for k, v in d.items():
        if isinstance(v, torch.Tensor):
            out[k] = v.detach().cpu().numpy()
        else:
            out[k] = np.asarray(v)
    return out


def golden_pvr():
    import iit.model_pairs as mp
    import torchvision
    from iit.tasks.mnist_pvr.pvr_hl import MNIST_PVR_HL, get_corr
    from iit.tasks.mnist_pvr.pvr_check_leaky_hl import MNIST_PVR_Leaky_HL, get_corr as get_corr_leaky
    from iit.utils.wrapper import HookedModuleWrapper, get_hook_points
    from iit.model_pairs.nodes import HLNode

    # B=64: with batch-statistic BatchNorm + ReLU + Adam a single fp32 rounding flip of one ReLU mask moves a
    # whole channel's gradient by O(1/rows); at B=8 that alone exceeds the 1e-3 loss bar after one Adam update
    B, size = 64, 56
    hook = "mod.layer3.mod.1.mod.conv2.hook_point"  # task_loader.py:41
    res = {"B": B}
    for mode in ("q", "c"):
        torch.manual_seed(0)
        net = torchvision.models.resnet18()
        net.fc = torch.nn.Linear(512, 10)
        ll = HookedModuleWrapper(net, name="resnet18", recursive=True, hook_self=False)
        hl = MNIST_PVR_HL(device="cpu")
        corr = get_corr(mode, hook, ll, (1, 3, size, size))
        # NB get_corr (pvr_hl.py:70) runs a zeros batch through the model in TRAIN mode, which moves
        # every BN running_var to 0.9; load the pinned weights afterwards so the fixture starts clean.
        ll.load_state_dict(restate.resnet18_state_dict(0))
        pair = mp.IITBehaviorModelPair(hl, ll, corr, training_args={"lr": 1e-3})
        opt = torch.optim.Adam(ll.parameters(), lr=1e-3)
        if mode == "q":
            res["hook_points"] = np.array(list(ll.hook_dict.keys()))
            res["conv_hook_points"] = np.array(get_hook_points(ll))
            res["state_dict_keys"] = np.array(list(ll.state_dict().keys()))
        nodes, iit, beh = [], [], []
        ll.train()
        for step in range(3):
            base = restate.synthetic_pvr(B, 100 + step, size)
            src = restate.synthetic_pvr(B, 200 + step, size)
            if step == 0:
                # do_intervention on its own (train mode), every HL node, before any update
                sd0 = {k: v.clone() for k, v in ll.state_dict().items()}
                for k in corr.keys():
                    hl_o, ll_o = pair.do_intervention(base, src, k)
                    res[f"{mode}/di/{k.name}/hl"] = hl_o
                    res[f"{mode}/di/{k.name}/ll"] = ll_o
                    if mode == "q" and k.name == "hook_tr":
                        res["q/di/hook_tr/cache_hook"] = pair.ll_cache[hook][:4]
                ll.load_state_dict(sd0)  # undo the BN running-stat updates of the probe runs
            before = pair.rng.bit_generator.state["state"]["state"]
            m = pair.run_train_step(base, src, pair.loss_fn, opt)
            del before
            iit.append(m["train/iit_loss"])
            beh.append(m["train/behavior_loss"])
        # replay node choices: same generator, same call order (base_model_pair.py:117-118)
        rng = np.random.default_rng(0)
        keys = list(corr.keys())
        nodes = [str(rng.choice(keys)) for _ in range(4)]
        res[f"{mode}/train/iit_loss"] = np.array(iit, dtype=np.float64)
        res[f"{mode}/train/behavior_loss"] = np.array(beh, dtype=np.float64)
        res[f"{mode}/nodes"] = np.array(nodes)
        ll.eval()
        with torch.no_grad():
            base = restate.synthetic_pvr(B, 300, size)
            src = restate.synthetic_pvr(B, 301, size)
            ev = pair.run_eval_step(base, src, pair.loss_fn)
        for k, v in ev.items():
            res[f"{mode}/eval/{k}"] = np.float64(v)
        sd = ll.state_dict()
        res[f"{mode}/final/fc_weight"] = sd["mod.fc.mod.weight"]
        res[f"{mode}/final/conv1_weight_sum"] = sd["mod.conv1.mod.weight"].double().sum()
        res[f"{mode}/final/l3_bn2_running_mean"] = sd["mod.layer3.mod.1.mod.bn2.mod.running_mean"]
        res[f"{mode}/final/l4_bn2_running_var"] = sd["mod.layer4.mod.1.mod.bn2.mod.running_var"]
        res[f"{mode}/final/num_batches_tracked"] = sd["mod.bn1.mod.num_batches_tracked"]

    # IIA sweep semantic (eval_causality.py:47-83 as intended: HLNode passed, leaky HL, eval mode)
    torch.manual_seed(0)
    net = torchvision.models.resnet18()
    net.fc = torch.nn.Linear(512, 10)
    ll = HookedModuleWrapper(net, name="resnet18", recursive=True, hook_self=False)
    ll.load_state_dict(restate.resnet18_state_dict(0))
    ll.eval()
    base = restate.synthetic_pvr(B, 400, size)
    for hp_name in ("mod.layer3.mod.1.mod.conv2.hook_point", "mod.layer1.mod.0.mod.conv1.hook_point"):
        hl = MNIST_PVR_Leaky_HL(device="cpu")
        corr = get_corr_leaky("q", hp_name, ll, (1, 3, size, size))
        pair = mp.IITBehaviorModelPair(hl, ll, corr)
        res[f"sweep/{hp_name}/nodes"] = np.array([k.name for k in corr.keys()])
        stats = []
        with torch.no_grad():
            for i, k in enumerate(corr.keys()):
                src = restate.synthetic_pvr_counterfactual(base, k.name, 500 + i)
                hl_o, ll_o = pair.do_intervention(base, src, k)
                changed = (src[1] != base[1]).float()
                top1 = torch.argmax(ll_o, dim=1)
                acc = (top1 == hl_o).float() * changed
                stats.append((acc.sum() / (changed.sum() + 1e-10)).item())
                if i == 0:
                    res[f"sweep/{hp_name}/first/hl"] = hl_o
                    res[f"sweep/{hp_name}/first/ll"] = ll_o
        res[f"sweep/{hp_name}/stats"] = np.array(stats, dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, "pvr.npz"), **to_np(res))
    print("wrote pvr.npz", len(res), "arrays")


def golden_ioi():
    import iit.model_pairs as mp
    from iit.tasks.ioi import corr, ioi_cfg, IOI_HL
    from iit.utils import node_picker
    import transformer_lens

    V, B = 101, 8
    cfg = dict(n_layers=ioi_cfg["n_layers"], n_heads=ioi_cfg["n_heads"], d_model=ioi_cfg["d_model"],
               d_head=ioi_cfg["d_head"], d_mlp=128, d_vocab=V, n_ctx=32, act_fn="gelu_new",
               normalization_type="LNPre", init_weights=True, device="cpu")
    ll = transformer_lens.HookedTransformer(cfg)
    sd = restate.transformer_state_dict(0, d_mlp=128, d_vocab=V, n_ctx=32)
    missing = ll.load_state_dict(sd, strict=False)
    assert not missing.unexpected_keys, missing
    assert all(k.endswith("mask") or k.endswith("IGNORE") for k in missing.missing_keys), missing
    (base0, names) = restate.synthetic_ioi(B, 100, d_vocab=V, n_names=20)
    hl = IOI_HL(d_vocab=V, names=names)
    targs = {"batch_size": B, "lr": 1e-3, "iit_weight": 1.0, "behavior_weight": 1.0, "strict_weight": 0.4,
             "next_token": False, "lr_scheduler": None, "clip_grad_norm": 1.0, "early_stop": True,
             "use_single_loss": False}
    pair = mp.IOI_ModelPair(ll_model=ll, hl_model=hl, corr=corr, training_args=targs)
    res = {"names": names}
    res["nodes_not_in_circuit"] = np.array([f"{n.name}|{n.index}" for n in pair.nodes_not_in_circuit])
    res["all_nodes"] = np.array([f"{n.name}|{n.index}" for n in node_picker.get_all_nodes(ll, corr.get_suffixes())])
    opt = torch.optim.Adam(ll.parameters(), lr=targs["lr"])
    ll.train()
    # do_intervention per HL node before any update
    base, _ = restate.synthetic_ioi(B, 100, d_vocab=V, n_names=20)
    src, _ = restate.synthetic_ioi(B, 200, d_vocab=V, n_names=20)
    for k in corr.keys():
        hl_o, ll_o = pair.do_intervention(base, src, k)
        res[f"di/{k.name}/hl_last"] = hl_o[:, -1, :]
        res[f"di/{k.name}/ll_last"] = ll_o[:, -1, :]
        res[f"di/{k.name}/iit_loss"] = pair.get_IIT_loss_over_batch(base, src, k, pair.loss_fn)
    hl_full = hl((base[0], None, None))
    res["hl/base_full"] = hl_full
    iit, beh, strict = [], [], []
    for step in range(3):
        base, _ = restate.synthetic_ioi(B, 100 + step, d_vocab=V, n_names=20)
        src, _ = restate.synthetic_ioi(B, 200 + step, d_vocab=V, n_names=20)
        m = pair.run_train_step(base, src, pair.loss_fn, opt)
        iit.append(m["train/iit_loss"]); beh.append(m["train/behavior_loss"]); strict.append(m["train/strict_loss"])
    res["train/iit_loss"] = np.array(iit, dtype=np.float64)
    res["train/behavior_loss"] = np.array(beh, dtype=np.float64)
    res["train/strict_loss"] = np.array(strict, dtype=np.float64)
    # replay of the choices: hl node then ll node per step, same stream (strict_iit_model_pair.py:50,62)
    rng = np.random.default_rng(0)
    keys = list(corr.keys())
    seq = []
    for _ in range(3):
        seq.append(str(rng.choice(keys)))
        n = rng.choice(pair.nodes_not_in_circuit)
        seq.append(f"{n.name}|{n.index}")
    res["choice_sequence"] = np.array(seq)
    ll.eval()
    with torch.no_grad():
        base, _ = restate.synthetic_ioi(B, 300, d_vocab=V, n_names=20)
        src, _ = restate.synthetic_ioi(B, 301, d_vocab=V, n_names=20)
        ev = pair.run_eval_step(base, src, pair.loss_fn)
    for k, v in ev.items():
        res[f"eval/{k}"] = np.asarray(v, dtype=np.float64)
    res["final/W_U_sum"] = ll.state_dict()["unembed.W_U"].double().sum()
    res["final/b3_W_in"] = ll.state_dict()["blocks.3.mlp.W_in"][:4, :8]
    np.savez_compressed(os.path.join(OUT, "ioi.npz"), **to_np(res))
    print("wrote ioi.npz", len(res), "arrays")


def golden_freeze_stopgrad():
    """FreezedModelPair / StopGradModelPair (iit/model_pairs/freeze_model_pair.py, stop_grad_pair.py), unmodified, on the
    small IOI fixture: three train steps and one eval step each, plus the node / parameter lists they build."""
    import transformer_lens
    from iit.model_pairs.freeze_model_pair import FreezedModelPair
    from iit.model_pairs.stop_grad_pair import StopGradModelPair
    from iit.tasks.ioi import corr, ioi_cfg, IOI_HL

    V, B = 101, 8
    cfg = dict(n_layers=ioi_cfg["n_layers"], n_heads=ioi_cfg["n_heads"], d_model=ioi_cfg["d_model"],
               d_head=ioi_cfg["d_head"], d_mlp=128, d_vocab=V, n_ctx=32, act_fn="gelu_new",
               normalization_type="LNPre", init_weights=True, device="cpu")
    res = {}
    for tag, cls in (("freeze", FreezedModelPair), ("stopgrad", StopGradModelPair)):
        ll = transformer_lens.HookedTransformer(cfg)
        ll.load_state_dict(restate.transformer_state_dict(0, d_mlp=128, d_vocab=V, n_ctx=32), strict=False)
        _, names = restate.synthetic_ioi(B, 100, d_vocab=V, n_names=20)
        hl = IOI_HL(d_vocab=V, names=names)
        pair = cls(hl, ll, corr, training_args={"lr": 1e-3, "lr_scheduler": None})
        res[f"{tag}/params_not_in_circuit"] = np.array([f"{p.name}|{p.index}" for p in pair.params_not_in_circuit])
        if tag == "stopgrad":
            res["stopgrad/nodes_not_in_circuit"] = np.array([f"{n.name}|{n.index}" for n in pair.ll_model.nodes_not_in_circuit])
            res["stopgrad/post_nodes_not_in_circuit"] = np.array([f"{n.name}|{n.index}" for n in pair.ll_model.post_nodes_not_in_circuit])
        opt = torch.optim.Adam(ll.parameters(), lr=1e-3)
        ll.train()
        iit, beh = [], []
        for step in range(3):
            base, _ = restate.synthetic_ioi(B, 100 + step, d_vocab=V, n_names=20)
            src, _ = restate.synthetic_ioi(B, 200 + step, d_vocab=V, n_names=20)
            m = pair.run_train_step(base, src, pair.loss_fn, opt)
            iit.append(m["train/iit_loss"]); beh.append(m["train/behavior_loss"])
            if step == 0:
                res[f"{tag}/after_step0/b2_W_in"] = ll.state_dict()["blocks.2.mlp.W_in"][:4, :8].clone()
                res[f"{tag}/after_step0/b0_W_in"] = ll.state_dict()["blocks.0.mlp.W_in"][:4, :8].clone()
        res[f"{tag}/train/iit_loss"] = np.array(iit, dtype=np.float64)
        res[f"{tag}/train/behavior_loss"] = np.array(beh, dtype=np.float64)
        ll.eval()
        with torch.no_grad():
            base, _ = restate.synthetic_ioi(B, 300, d_vocab=V, n_names=20)
            src, _ = restate.synthetic_ioi(B, 301, d_vocab=V, n_names=20)
            ev = pair.run_eval_step(base, src, pair.loss_fn)
        for k, v in ev.items():
            res[f"{tag}/eval/{k}"] = np.asarray(v, dtype=np.float64)
    np.savez_compressed(os.path.join(OUT, "ioi_freeze_stopgrad.npz"), **to_np(res))
    print("wrote ioi_freeze_stopgrad.npz", len(res), "arrays")


def _ioi_ablation_setup(bias):
    """Small IOI pair (as golden_ioi) on 48 prompts with 4 names; `bias` is added to one name's unembed bias so that the
    accuracy-based statistics are not trivially zero."""
    import iit.model_pairs as mp
    from iit.tasks.ioi import corr, IOI_HL
    import transformer_lens
    V, N = 101, 48
    cfg = dict(n_layers=6, n_heads=4, d_model=64, d_head=16, d_mlp=128, d_vocab=V, n_ctx=32, act_fn="gelu_new",
               normalization_type="LNPre", init_weights=True, device="cpu")
    ll = transformer_lens.HookedTransformer(cfg)
    sd, data, names = restate.ioi_ablation_fixture(bias)
    ll.load_state_dict(sd, strict=False)
    hl = IOI_HL(d_vocab=V, names=names)
    pair = mp.IOI_ModelPair(ll_model=ll, hl_model=hl, corr=corr, training_args={"next_token": False})
    ll.eval()
    return pair, data


def golden_ioi_ablations():
    """iit/utils/eval_ablations.py run unmodified: resample (KL), mean and zero ablation."""
    from iit.utils import eval_ablations as ea
    from iit.utils.eval_datasets import IITUniqueDataset
    from iit.utils.iit_dataset import IITDataset
    from iit.model_pairs.nodes import LLNode
    from iit.utils.index import Ix

    def key(n):
        return f"{n.name}|{n.index}"

    res = {}
    pair, data = _ioi_ablation_setup(0.0)
    ds = IITDataset(data, data, seed=0, device="cpu")
    for nt in ("n", "c"):
        torch.manual_seed(0)
        with torch.no_grad():
            r = ea.check_causal_effect(pair, ds, batch_size=16, node_type=nt)
        res[f"resample_kl/{nt}/nodes"] = np.array([key(n) for n in r])
        res[f"resample_kl/{nt}/values"] = np.array(list(r.values()), dtype=np.float64)
    uni = IITUniqueDataset(data, data, seed=0, device="cpu")
    torch.manual_seed(0)
    with torch.no_grad():
        mc = ea.get_mean_cache(pair, uni, batch_size=16)
    for name in ("blocks.1.attn.hook_z", "blocks.2.mlp.hook_post", "blocks.0.hook_resid_pre"):
        res[f"mean_cache/{name}"] = mc[name]
    # logits of the base run with one node mean- / zero-ablated, first 8 prompts
    x8 = torch.stack([d[0] for d in data[:8]])
    for node in (LLNode("blocks.1.attn.hook_z", Ix[:, :, 2, :]), LLNode("blocks.3.mlp.hook_post", Ix[[None]])):
        for use_mean in (True, False):
            hook = ea.make_ablation_hook(node, mc, use_mean)
            with torch.no_grad():
                out = pair.ll_model.run_with_hooks(x8, fwd_hooks=[(node.name, hook)])
            res[f"ablated_logits/{key(node)}/{'mean' if use_mean else 'zero'}"] = out[:, -1, :]
    # the statistics themselves where every decision has a wide margin (one name's bias + 4)
    pair4, data4 = _ioi_ablation_setup(4.0)
    uni4 = IITUniqueDataset(data4, data4, seed=0, device="cpu")
    for use_mean in (True, False):
        torch.manual_seed(0)
        with torch.no_grad():
            mc4 = ea.get_mean_cache(pair4, uni4, batch_size=16) if use_mean else None
            a = ea.check_causal_effect_on_ablation(pair4, uni4, batch_size=16, node_type="n", mean_cache=mc4)
            b = ea.check_causal_effect_on_ablation(pair4, uni4, batch_size=16, node_type="c", mean_cache=mc4)
        tag = "mean" if use_mean else "zero"
        res[f"ablate/{tag}/n/nodes"] = np.array([key(n) for n in a])
        res[f"ablate/{tag}/n/values"] = np.array(list(a.values()), dtype=np.float64)
        res[f"ablate/{tag}/c/nodes"] = np.array([key(n) for n in b])
        res[f"ablate/{tag}/c/values"] = np.array(list(b.values()), dtype=np.float64)
    with torch.no_grad():
        base_ll = pair4.ll_model(torch.stack([d[0] for d in data4]))[:, -1, :]
        base_hl = pair4.hl_model((torch.stack([d[0] for d in data4]), None, None))[:, -1, :]
    res["ablate/base_accuracy"] = (base_ll.argmax(-1) == base_hl.argmax(-1)).float().mean()
    np.savez_compressed(os.path.join(OUT, "ioi_ablations.npz"), **to_np(res))
    print("wrote ioi_ablations.npz", len(res), "arrays")


if __name__ == "__main__":
    ref_env.setup()
    import iit.model_pairs  # noqa: F401  (import first: the reference has an import cycle otherwise)
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(4)
    only = sys.argv[1:]
    if not only or "pvr" in only:
        golden_pvr()
    if not only or "ioi" in only:
        golden_ioi()
    if not only or "ablations" in only:
        golden_ioi_ablations()
    if not only or "freeze" in only:
        golden_freeze_stopgrad()
