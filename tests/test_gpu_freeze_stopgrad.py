"""GPU: FreezedModelPair / StopGradModelPair through the C-ABI against
  (1) tests/golden/ioi_freeze_stopgrad.npz -- the UNMODIFIED reference classes run through the transformer_lens shim, and
  (2) oracle/restate.py:OracleFreezePair on the CPU, term by term from synchronised weights.
Bars: losses 1e-3 relative (fp32), gradients 1e-2 in relative L2, eval counts equal."""
import os

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu

from oracle import restate as R  # noqa: E402

GOLD = np.load(os.path.join(os.path.dirname(__file__), "golden", "ioi_freeze_stopgrad.npz"))
DEV = "cuda"
V, DM, B = 101, 128, 8


def dev(inp):
    return tuple(t.to(DEV) for t in inp)


def rel(a, b):
    return abs(a - b) / max(abs(b), 1e-12)


def batch(seed):
    return R.synthetic_ioi(B, seed, d_vocab=V, n_names=20)[0]


def make(cls_name):
    import iit_b200.model_pairs as mp
    from iit_b200.tasks.ioi import IOI_HL, corr, make_ll_model
    ll = make_ll_model(device=DEV, d_mlp=DM, d_vocab=V, n_ctx=32)
    ll.load_state_dict(R.transformer_state_dict(0, d_mlp=DM, d_vocab=V, n_ctx=32))
    _, names = R.synthetic_ioi(B, 100, d_vocab=V, n_names=20)
    hl = IOI_HL(V, names).to(DEV)
    return getattr(mp, cls_name)(hl, ll, corr, training_args={"lr": 1e-3, "lr_scheduler": None}), names


@pytest.mark.parametrize("tag,cls_name", [("freeze", "FreezedModelPair"), ("stopgrad", "StopGradModelPair")])
def test_free_running_steps_match_the_reference_golden(tag, cls_name):
    pair, _ = make(cls_name)
    assert [f"{p.name}|{p.index}" for p in pair.params_not_in_circuit] == GOLD[f"{tag}/params_not_in_circuit"].tolist()
    opt = pair.make_optimizer()
    pair.ll_model.train()
    for step in range(3):
        m = pair.run_train_step(dev(batch(100 + step)), dev(batch(200 + step)), pair.loss_fn, opt)
        tol = 1e-3 if step == 0 else 1e-2        # free-running after Adam updates: see tests/test_gpu_pvr_parity.py
        assert rel(m["train/iit_loss"], GOLD[f"{tag}/train/iit_loss"][step]) < tol, (step, m)
        assert rel(m["train/behavior_loss"], GOLD[f"{tag}/train/behavior_loss"][step]) < tol, (step, m)
        if step == 0:
            sd = pair.ll_model.state_dict()
            for key, name in (("b2_W_in", "blocks.2.mlp.W_in"), ("b0_W_in", "blocks.0.mlp.W_in")):
                np.testing.assert_allclose(sd[name][:4, :8].cpu().numpy(), GOLD[f"{tag}/after_step0/{key}"], rtol=0, atol=3e-4)
    pair.ll_model.eval()
    with torch.no_grad():
        ev = pair.run_eval_step(dev(batch(300)), dev(batch(301)), pair.loss_fn)
    assert rel(ev["val/iit_loss"], float(GOLD[f"{tag}/eval/val/iit_loss"])) < 1e-2
    assert ev["val/IIA"] == float(GOLD[f"{tag}/eval/val/IIA"]) and ev["val/accuracy"] == float(GOLD[f"{tag}/eval/val/accuracy"])


@pytest.mark.parametrize("stop", [False, True])
def test_terms_from_synchronised_weights_match_the_restatement(stop):
    """IIT term, behaviour term, then the NEXT IIT term (whose source run still carries the stop-gradient hooks the
    behaviour forward left attached) against OracleFreezePair, weights re-synchronised before each."""
    pair, names = make("StopGradModelPair" if stop else "FreezedModelPair")
    ll = pair.ll_model.model if stop else pair.ll_model
    oll = R.TransformerOracle(R.transformer_state_dict(0, d_mlp=DM, d_vocab=V, n_ctx=32), 6, 4, 16)
    opair = R.OracleFreezePair(oll, R.ioi_corr(), names, V, stop)
    opt = pair.make_optimizer()
    ll.train()
    b, s = batch(100), batch(200)
    bd, sd_ = dev(b), dev(s)
    node = pair.sample_hl_name()
    terms = [("iit", lambda: pair.get_IIT_loss_over_batch(bd, sd_, node, pair.loss_fn), lambda: opair.iit_loss(b, s, str(node))),
             ("behaviour", lambda: pair.get_behaviour_loss_over_batch(bd, pair.loss_fn), lambda: opair.behaviour_loss(b)),
             ("iit after behaviour", lambda: pair.get_IIT_loss_over_batch(bd, sd_, node, pair.loss_fn),
              lambda: opair.iit_loss(b, s, str(node)))]
    for name, mine, theirs in terms:
        ll.load_state_dict({k: v.detach() for k, v in oll.p.items()})
        opt.zero_grad()
        lm = mine()
        lm.backward()
        opair.opt.zero_grad()
        lo = theirs()
        lo.backward()
        assert rel(lm.item(), lo.item()) < 1e-3, (name, lm.item(), lo.item())
        num = den = 0.0
        for k, p in ll.named_parameters():
            og = oll.p[k].grad
            og = torch.zeros_like(oll.p[k]).double() if og is None else og.double()
            g = p.grad.detach().cpu().double()
            num += float(((g - og) ** 2).sum())
            den += float((og ** 2).sum())
        assert (num / den) ** 0.5 < 1e-2, (name, (num / den) ** 0.5)
        if stop and name == "behaviour":
            named = dict(ll.named_parameters())
            assert float(named["blocks.3.mlp.W_in"].grad.abs().max()) == 0.0      # outside the circuit: no gradient
            assert float(named["blocks.0.mlp.W_in"].grad.abs().max()) > 0.0
        torch.nn.utils.clip_grad_norm_(oll.parameters(), 1.0)
        opair.opt.step()
    if stop:   # the leftover hooks changed the second source run: its cache differs from a clean source run's
        assert not any(hp.has_hooks() for hp in ll.hook_dict.values())


def test_div_scalar_and_3d_cross_entropy_kernels():
    from iit_b200 import ops
    g = torch.Generator().manual_seed(0)
    x = torch.randn(5, 16, 4, 16, generator=g)
    xd = x.to(DEV).requires_grad_(True)
    y = ops.div_scalar(xd, 1e6)
    assert torch.equal(y.cpu(), x / 1e6)
    y.backward(torch.ones_like(y))
    assert torch.equal(xd.grad.cpu(), torch.ones_like(x) / 1e6)
    # CrossEntropyLoss on [N, C, d1] (class dimension 1) with probability and with index targets
    lg = torch.randn(6, 16, 37, generator=g).requires_grad_(True)
    tp = torch.randn(6, 16, 37, generator=g) * 3           # the IOI HL "logits as probabilities" case: any real numbers
    want = torch.nn.functional.cross_entropy(lg, tp)
    want.backward()
    ld = lg.detach().to(DEV).requires_grad_(True)
    got = ops.cross_entropy(ld, tp.to(DEV))
    got.backward()
    assert rel(got.item(), want.item()) < 1e-5
    torch.testing.assert_close(ld.grad.cpu(), lg.grad, rtol=1e-4, atol=1e-6)
    ti = torch.randint(0, 16, (6, 37), generator=g)
    want_i = torch.nn.functional.cross_entropy(lg.detach(), ti)
    assert rel(ops.cross_entropy(lg.detach().to(DEV), ti.to(DEV)).item(), want_i.item()) < 1e-5
