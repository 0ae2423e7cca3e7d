"""What bf16 ACTIVATION STORAGE alone does to the gradient of the randomly initialised ResNet-18, measured on the CPU
oracle (no kernels involved): the fp32 restatement of the reference (oracle/restate.py) is run once as is and once with
every hook-point tensor rounded to bfloat16 on the way forward (and, separately, every gradient on the way back).

This is the yardstick for the bf16 gradient bars in tests/test_gpu_contract_configs.py and tests/test_gpu_hook_sites.py:
the reference has no bf16 path, so "bf16 parity" is stated against the fp32 oracle, and this test shows how far rounding
by itself moves the gradient (tens of percent in relative L2 at initialisation, growing towards the stem, almost all of it
from the FORWARD rounding) while the loss moves by ~1e-3."""
import torch
import torch.nn.functional as F

from oracle import restate as R


class _RoundFwd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return x.bfloat16().float()

    @staticmethod
    def backward(ctx, g):
        return g


class _RoundBwd(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return x.clone()

    @staticmethod
    def backward(ctx, g):
        return g.bfloat16().float()


def _grads(mode, batch):
    oll = R.ResNet18Oracle(R.resnet18_state_dict(0))
    x, y, _ = R.synthetic_pvr(batch, 11)
    names = {}
    with torch.no_grad():
        R.ResNet18Oracle(R.resnet18_state_dict(0))(x[:2], cache=names)
    fn = {"fwd": _RoundFwd.apply, "bwd": _RoundBwd.apply}.get(mode)
    hooks = {n: fn for n in names if n not in ("mod.fc.hook_point", "mod.avgpool.hook_point")} if fn else {}
    loss = F.cross_entropy(oll(x, hooks=hooks), y)
    loss.backward()
    return {k: p.grad.double() for k, p in oll.p.items()}, loss.item()


def _rel_l2(g, g0):
    num = sum(float((g[k] - g0[k]).norm() ** 2) for k in g0)
    den = sum(float(g0[k].norm() ** 2) for k in g0)
    return (num / den) ** 0.5


def test_forward_rounding_to_bf16_moves_the_fp32_gradient_by_tens_of_percent():
    torch.set_num_threads(max(1, torch.get_num_threads()))
    g0, l0 = _grads("none", 64)
    gf, lf = _grads("fwd", 64)
    gb, lb = _grads("bwd", 64)
    e_fwd, e_bwd = _rel_l2(gf, g0), _rel_l2(gb, g0)
    print(f"bf16 rounding model at batch 64: loss {l0:.6f} -> {lf:.6f}; gradient rel-L2 forward-rounding {e_fwd:.3f}, "
          f"backward-rounding {e_bwd:.3f}")
    assert abs(lf - l0) / l0 < 5e-3 and lb == l0
    assert e_fwd > 0.15            # forward rounding alone: far above any kernel-error bar one could set
    assert e_bwd < 0.05            # rounding the gradients themselves is benign
    # and it accumulates from the head towards the stem, as the kernels' error does
    first = float((gf["mod.conv1.mod.weight"] - g0["mod.conv1.mod.weight"]).norm() / g0["mod.conv1.mod.weight"].norm())
    last = float((gf["mod.fc.mod.weight"] - g0["mod.fc.mod.weight"]).norm() / g0["mod.fc.mod.weight"].norm())
    assert first > 5 * last
