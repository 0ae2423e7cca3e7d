"""GPU parity of each kernel, called through the C-ABI (iit_b200.ops -> libiit_b200.so), against
the oracle (oracle/restate.py) for the integer / index work -- bit-exact -- and against a plain
PyTorch fp32 CPU evaluation of the same op for the floating-point kernels (tolerances stated inline)."""
import math

import numpy as np
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu

from oracle import restate as R  # noqa: E402


@pytest.fixture(scope="module")
def ops():
    from iit_b200 import ops
    return ops


@pytest.fixture(scope="module")
def Ix():
    from iit_b200.utils.index import Ix
    return Ix


DEV = "cuda"


# ------------------------------------------------------------------------------------------ patch
def _patch_cases(Ix):
    return [
        ((8, 256, 4, 4), Ix[None, None, :2, :2], "q tl"),
        ((8, 256, 4, 4), Ix[None, None, 2:4, 2:4], "q br"),
        ((8, 256, 6, 6), Ix[None, None, :3, 3:6], "q 84px"),
        ((8, 256, 4, 4), Ix[None, 64:128, None, None], "c mode"),
        ((8, 256, 4, 4), Ix[[None]], "whole"),
        ((4, 16, 4, 16), Ix[:, :, 2, :], "one head"),
        ((4, 16, 4, 16), Ix[:, -1], "negative int"),
        ((4, 16, 3072), Ix[[None]], "mlp whole"),
        ((5, 7, 3), Ix[1:4, :, 2], "odd sizes"),
        ((3, 5), Ix[:, 2:2], "empty box"),
        ((6,), Ix[2:5], "rank 1"),
    ]


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.int64])
@pytest.mark.parametrize("layout", ["contiguous", "channels_last", "mixed"])
def test_patch_bit_exact(ops, Ix, dtype, layout):
    g = torch.Generator().manual_seed(0)
    for shape, idx, name in _patch_cases(Ix):
        def mk():
            v = torch.randn(shape, generator=g) * 100
            return v.to(dtype)
        act, src = mk(), mk()
        want = R.patch(act, src, idx.as_index)           # reference semantics on CPU
        a, s = act.to(DEV), src.to(DEV)
        if len(shape) == 4 and layout != "contiguous":
            a = a.contiguous(memory_format=torch.channels_last)
            if layout == "channels_last":
                s = s.contiguous(memory_format=torch.channels_last)
        out = ops.patch(a, s, idx, inplace=True)
        assert out.data_ptr() == a.data_ptr(), name
        assert torch.equal(out.cpu(), want), f"{name} {dtype} {layout}"
        cap = ops.capture_box(s, idx)
        assert torch.equal(cap.cpu().reshape(-1), src[idx.as_index].reshape(-1)), f"capture {name}"


def test_patch_list_index_falls_back(ops, Ix):
    act, src = torch.randn(4, 6, 5), torch.randn(4, 6, 5)
    idx = Ix[:, [0, 3], :]
    out = ops.patch(act.to(DEV), src.to(DEV), idx, inplace=False)
    assert torch.equal(out.cpu(), R.patch(act, src, idx.as_index))


def test_patch_gradient_is_copyslices(ops, Ix):
    for shape, idx, name in _patch_cases(Ix):
        x = torch.randn(shape, requires_grad=True)
        src = torch.randn(shape)
        w = torch.randn(shape)
        (R.patch(x * 2, src, idx.as_index) * w).sum().backward()
        xg = x.detach().to(DEV).requires_grad_(True)
        out = ops.patch(xg * 2, src.to(DEV), idx, inplace=True)
        (out * w.to(DEV)).sum().backward()
        got = torch.zeros(shape) if xg.grad is None else xg.grad.cpu()  # whole-tensor patch: no gradient at all
        assert torch.equal(got, x.grad), name


def test_patch_rejects_cpu_tensors(ops, Ix):
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        ops.patch_(torch.zeros(4, 4), torch.zeros(4, 4), Ix[[None]])


# --------------------------------------------------------------------------------------------- HL
def test_hl_pvr_bit_exact(ops):
    g = torch.Generator().manual_seed(1)
    for batch in (1, 7, 4096):
        bv = torch.randint(0, 10, (batch, 4), generator=g)
        sv = torch.randint(0, 10, (batch, 4), generator=g)
        assert torch.equal(ops.hl_pvr(bv.to(DEV), None, -1).cpu(), R.pvr_hl_forward(bv))
        for i, h in enumerate(R.PVR_HOOKS):
            got = ops.hl_pvr(bv.to(DEV), sv.to(DEV), i).cpu()
            assert torch.equal(got, R.pvr_hl_intervened(bv, sv, h)), h


def _rand_tokens(batch, seq, vocab, names, g):
    tok = torch.randint(0, vocab, (batch, seq), generator=g)
    # plenty of names and duplicates
    m = torch.rand(batch, seq, generator=g) < 0.5
    ok = names[names < vocab]
    tok[m] = ok[torch.randint(0, len(ok), (int(m.sum()),), generator=g)]
    return tok


@pytest.mark.parametrize("vocab,seq", [(21, 9), (101, 16), (50257, 16)])
def test_hl_ioi_bit_exact(ops, vocab, seq):
    g = torch.Generator().manual_seed(2)
    names = torch.tensor([10, 20, 30]) if vocab == 21 else torch.randperm(vocab - 1, generator=g)[:20].sort().values
    batch = 64 if vocab < 1000 else 8
    base = _rand_tokens(batch, seq, vocab, names, g)
    src = _rand_tokens(batch, seq, vocab, names, g)
    # reference's own vector (iit/tasks/ioi/test_ioi.py:34-45)
    if vocab == 21:
        base[0] = torch.tensor([3, 10, 4, 10, 5, 9, 2, 6, 5])
    nd, bd, sd = names.to(DEV), base.to(DEV), src.to(DEV)
    dense = ops.hl_ioi_dense(bd, None, None, nd, vocab).cpu()
    want = R.ioi_hl_forward(base, names, vocab)
    assert torch.equal(dense, want)
    assert torch.equal(ops.hl_ioi_last_argmax(bd, None, None, nd, vocab).cpu(), torch.argmax(want[:, -1, :], dim=-1))
    for node in R.IOI_HL_HOOKS:
        want = R.ioi_hl_intervened(base, src, names, vocab, node)
        assert torch.equal(ops.hl_ioi_dense(bd, sd, node, nd, vocab).cpu(), want), node
        assert torch.equal(ops.hl_ioi_last_argmax(bd, sd, node, nd, vocab).cpu(),
                           torch.argmax(want[:, -1, :], dim=-1)), node


# ------------------------------------------------------------------------------------- CE and IIA
@pytest.mark.parametrize("rows,classes", [(8, 10), (4096, 10), (5, 101), (6, 50257)])
def test_ce_index_and_prob(ops, rows, classes):
    g = torch.Generator().manual_seed(3)
    logits = (torch.randn(rows, classes, generator=g) * 3).requires_grad_(True)
    tgt = torch.randint(0, classes, (rows,), generator=g)
    want = F.cross_entropy(logits, tgt)
    (want * 0.4).backward()
    lg = logits.detach().to(DEV).requires_grad_(True)
    got = ops.cross_entropy(lg, tgt.to(DEV))
    (got * 0.4).backward()
    # fp32: relative 1e-5 on the loss, 1e-5 absolute on gradients of magnitude <= 1/rows
    assert math.isclose(got.item(), want.item(), rel_tol=1e-5)
    torch.testing.assert_close(lg.grad.cpu(), logits.grad, rtol=1e-4, atol=1e-7)
    # probability target: one-hot must equal the index form; soft targets must match torch
    onehot = F.one_hot(tgt, classes).float()
    soft = torch.softmax(torch.randn(rows, classes, generator=g), -1)
    for prob in (onehot, soft):
        l2 = logits.detach().clone().requires_grad_(True)
        w2 = F.cross_entropy(l2, prob)
        w2.backward()
        lg2 = logits.detach().to(DEV).requires_grad_(True)
        g2 = ops.cross_entropy(lg2, prob.to(DEV))
        g2.backward()
        assert math.isclose(g2.item(), w2.item(), rel_tol=1e-5)
        torch.testing.assert_close(lg2.grad.cpu(), l2.grad, rtol=1e-4, atol=1e-7)


def test_ce_strided_rows(ops):
    """last-position rows of a [B,S,V] tensor are read in place (row stride S*V)."""
    g = torch.Generator().manual_seed(4)
    full = torch.randn(6, 16, 101, generator=g)
    tgt = torch.randint(0, 101, (6,), generator=g)
    want = F.cross_entropy(full[:, -1, :], tgt)
    got = ops.cross_entropy(full.to(DEV)[:, -1, :], tgt.to(DEV))
    assert math.isclose(got.item(), want.item(), rel_tol=1e-5)


@pytest.mark.parametrize("rows,classes", [(9, 10), (4096, 10), (7, 1025), (5, 50257)])
def test_iia_count_bit_exact(ops, rows, classes):
    g = torch.Generator().manual_seed(5)
    logits = torch.randn(rows, classes, generator=g)
    logits[0, :] = 0.0                       # all tied -> first index
    logits[1, -1] = logits[1, 3] = 50.0      # two-way tie -> lower index
    if rows > 2:
        logits[2, 5] = float("nan")          # NaN is maximal in torch.argmax
    hl = torch.randint(0, classes, (rows,), generator=g)
    hl[::2] = torch.argmax(logits, -1)[::2]
    changed = torch.rand(rows, generator=g) < 0.6
    top1_want = torch.argmax(logits, dim=-1)
    counts, top1 = ops.iia_count(logits.to(DEV), hl.to(DEV), want_top1=True)
    assert torch.equal(top1.cpu(), top1_want)
    assert counts.tolist() == [int((top1_want == hl).sum()), rows]
    counts, _ = ops.iia_count(logits.to(DEV), hl.to(DEV), changed=changed.to(DEV))
    assert counts.tolist() == [int(((top1_want == hl) & changed).sum()), int(changed.sum())]
    # accumulation across calls
    counts, _ = ops.iia_count(logits.to(DEV), hl.to(DEV), counts=counts)
    assert counts.tolist() == [int(((top1_want == hl) & changed).sum()) + int((top1_want == hl).sum()),
                               int(changed.sum()) + rows]
    # bf16 logits
    lb = logits.to(torch.bfloat16)
    _, t2 = ops.iia_count(lb.to(DEV), None, want_top1=True)
    assert torch.equal(t2.cpu(), torch.argmax(lb.float(), dim=-1))


# ------------------------------------------------------------------------------- ResNet layers
RESNET_CONVS = [  # (cin, cout, k, stride, pad, hw)
    (3, 64, 7, 2, 3, 56), (64, 64, 3, 1, 1, 14), (64, 128, 3, 2, 1, 14), (64, 128, 1, 2, 0, 14),
    (128, 128, 3, 1, 1, 7), (128, 256, 3, 2, 1, 7), (256, 256, 3, 1, 1, 4), (256, 512, 3, 2, 1, 4),
    (512, 512, 3, 1, 1, 2), (256, 512, 1, 2, 0, 4), (64, 64, 3, 1, 1, 21),
]


@pytest.mark.parametrize("cfg", RESNET_CONVS)
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_conv_fprop_dgrad_wgrad(ops, cfg, dtype):
    from iit_b200.nn.resnet import Conv2d
    cin, cout, k, stride, pad, hw = cfg
    g = torch.Generator().manual_seed(6)
    n = 5
    x = torch.randn(n, cin, hw, hw, generator=g)
    w = torch.randn(cout, cin, k, k, generator=g) / math.sqrt(cin * k * k)
    if dtype == torch.bfloat16:  # compare against fp32 math on the bf16-rounded operands
        x, w = x.bfloat16().float(), w.bfloat16().float()
    x.requires_grad_(True)
    w.requires_grad_(True)
    y = F.conv2d(x, w, None, stride, pad)
    gy = torch.randn(y.shape, generator=g)
    if dtype == torch.bfloat16:
        gy = gy.bfloat16().float()
    y.backward(gy)
    conv = Conv2d(cin, cout, k, stride, pad).to(DEV)
    with torch.no_grad():
        conv.weight.copy_(w)
    xd = x.detach().to(DEV).to(dtype).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    sums = ops.new_bn_sums(cout, DEV)
    yd = conv(xd, sums)
    assert yd.shape == y.shape and yd.is_contiguous(memory_format=torch.channels_last)
    yd.backward(gy.to(DEV).to(dtype))
    # fp32 path: FFMA accumulation order differs from the CPU's -> 2e-4 of the output scale.
    # bf16 path: operands identical, outputs rounded to bf16 -> 1e-2 of the output scale.
    tol = 2e-4 if dtype == torch.float32 else 1e-2
    def close(a, b, what):
        scale = b.abs().max().item() + 1e-12
        err = (a.float().cpu() - b).abs().max().item()
        assert err <= tol * scale, f"{what}: err {err:.3e} scale {scale:.3e} cfg {cfg} {dtype}"
    close(yd, y.detach(), "fprop")
    close(xd.grad, x.grad, "dgrad")
    close(conv.weight.grad, w.grad, "wgrad")
    # fused BN statistics of the stored output
    ys = yd.detach().float().cpu()
    rows = ys.numel() // cout
    torch.testing.assert_close(sums[0].cpu().float() / rows, ys.mean(dim=(0, 2, 3)), rtol=1e-3, atol=1e-4)
    torch.testing.assert_close(sums[1].cpu().float() / rows, (ys * ys).mean(dim=(0, 2, 3)), rtol=1e-3, atol=1e-4)


@pytest.mark.parametrize("c,hw", [(64, 28), (128, 7), (512, 2)])
@pytest.mark.parametrize("relu,res", [(False, False), (True, False), (True, True)])
@pytest.mark.parametrize("training", [True, False])
def test_bn_act_fwd_bwd(ops, c, hw, relu, res, training):
    from iit_b200.nn.resnet import BatchNorm2d
    g = torch.Generator().manual_seed(7)
    n = 6
    x = (torch.randn(n, c, hw, hw, generator=g) * 2 + 0.5).requires_grad_(True)
    r = torch.randn(n, c, hw, hw, generator=g).requires_grad_(True) if res else None
    gamma = (torch.rand(c, generator=g) + 0.5).requires_grad_(True)
    beta = torch.randn(c, generator=g).requires_grad_(True)
    rm, rv = torch.randn(c, generator=g), torch.rand(c, generator=g) + 0.5
    rm0, rv0 = rm.clone(), rv.clone()
    y = F.batch_norm(x, rm, rv, gamma, beta, training, 0.1, 1e-5)
    if res:
        y = y + r
    if relu:
        y = F.relu(y)
    gy = torch.randn(y.shape, generator=g)
    y.backward(gy)
    bn = BatchNorm2d(c).to(DEV)
    bn.train(training)
    with torch.no_grad():
        bn.weight.copy_(gamma); bn.bias.copy_(beta); bn.running_mean.copy_(rm0); bn.running_var.copy_(rv0)
    xd = x.detach().to(DEV).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    rd = r.detach().to(DEV).contiguous(memory_format=torch.channels_last).requires_grad_(True) if res else None
    yd = bn(xd, residual=rd, relu=relu)
    yd.backward(gy.to(DEV))
    # fp32 elementwise + reductions: 1e-4 relative / 1e-5 absolute
    kw = dict(rtol=1e-4, atol=2e-5)
    torch.testing.assert_close(yd.cpu(), y.detach(), **kw)
    torch.testing.assert_close(xd.grad.cpu(), x.grad, **kw)
    torch.testing.assert_close(bn.weight.grad.cpu(), gamma.grad, rtol=1e-4, atol=1e-4)
    torch.testing.assert_close(bn.bias.grad.cpu(), beta.grad, rtol=1e-4, atol=1e-4)
    if res:
        torch.testing.assert_close(rd.grad.cpu(), r.grad, **kw)
    torch.testing.assert_close(bn.running_mean.cpu(), rm, **kw)
    torch.testing.assert_close(bn.running_var.cpu(), rv, **kw)
    assert int(bn.num_batches_tracked) == (1 if training else 0)


@pytest.mark.parametrize("c,hw", [(64, 28), (64, 13), (128, 6)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_bn_relu_maxpool_fused_equals_unfused_chain(ops, c, hw, dtype):
    """The fused stem tail (iit_bn_relu_maxpool_fwd/bwd) against (i) the unfused kernels bn_act(relu) -> maxpool:
    forward values and window winners bit-identical, gradients to fp32 rounding; (ii) plain PyTorch fp32 on the CPU
    (fp32 case: 1e-4 relative; the bf16 case is covered by (i))."""
    g = torch.Generator().manual_seed(11)
    n = 5
    x = (torch.randn(n, c, hw, hw, generator=g) * 2 + 0.3)
    gamma, beta = torch.rand(c, generator=g) + 0.5, torch.randn(c, generator=g) * 0.3
    rm, rv = torch.zeros(c), torch.ones(c)

    def run(fused):
        xd = x.to(DEV).to(dtype).contiguous(memory_format=torch.channels_last).requires_grad_(True)
        gd, bd = gamma.to(DEV).requires_grad_(True), beta.to(DEV).requires_grad_(True)
        rmd, rvd = rm.to(DEV), rv.to(DEV)
        if fused:
            y = ops.bn_relu_maxpool(xd, gd, bd, rmd, rvd, True, 0.1, 1e-5)
        else:
            y = ops.maxpool3x3s2(ops.bn_act(xd, gd, bd, None, rmd, rvd, True, 0.1, 1e-5, True))
        gy = torch.randn(y.shape, generator=torch.Generator().manual_seed(12)).to(DEV).to(dtype)
        y.backward(gy)
        return y.detach(), xd.grad, gd.grad, bd.grad, rmd, rvd, gy

    yf, dxf, dgf, dbf, rmf, rvf, gy = run(True)
    yu, dxu, dgu, dbu, rmu, rvu, _ = run(False)
    assert torch.equal(yf, yu)
    with torch.no_grad():     # no backward to follow: the winners' raw values are not written (NULL xwin)
        xd = x.to(DEV).to(dtype).contiguous(memory_format=torch.channels_last)
        yn = ops.bn_relu_maxpool(xd, gamma.to(DEV), beta.to(DEV), rm.to(DEV), rv.to(DEV), True, 0.1, 1e-5)
    assert torch.equal(yn, yf)
    assert torch.equal(rmf, rmu) and torch.equal(rvf, rvu)
    if dtype == torch.float32:
        torch.testing.assert_close(dxf, dxu, rtol=1e-4, atol=1e-5)
        torch.testing.assert_close(dgf, dgu, rtol=1e-3, atol=1e-3)
        torch.testing.assert_close(dbf, dbu, rtol=1e-3, atol=1e-3)
    else:
        # bf16: activations that tie after rounding send their gradient to the first of them in the unfused chain
        # and to the largest pre-rounding one here (as fp32 would): a few elements differ by a whole gradient value.
        # The unfused chain also rounds the pooled-gradient tensor to bf16 before the per-channel sums.
        bad = ((dxf.float() - dxu.float()).abs() > 2e-2 + 2e-2 * dxu.float().abs()).float().mean().item()
        assert bad < 0.02, bad
        torch.testing.assert_close(dgf, dgu, rtol=5e-2, atol=1.5)
        torch.testing.assert_close(dbf, dbu, rtol=5e-2, atol=1.5)
    if dtype == torch.float32:
        xr = x.clone().requires_grad_(True)
        gr, br = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
        yr = F.max_pool2d(F.relu(F.batch_norm(xr, rm.clone(), rv.clone(), gr, br, True, 0.1, 1e-5)), 3, 2, 1)
        yr.backward(gy.cpu().float())
        torch.testing.assert_close(yf.cpu(), yr.detach(), rtol=1e-4, atol=2e-5)
        torch.testing.assert_close(dxf.cpu(), xr.grad, rtol=1e-4, atol=2e-5)
        torch.testing.assert_close(dgf.cpu(), gr.grad, rtol=1e-4, atol=1e-4)
        torch.testing.assert_close(dbf.cpu(), br.grad, rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("c", [64, 256, 4])
def test_bn_bf16_fast_and_generic_paths_match_fp32_reference(ops, c):
    """bf16 BatchNorm(+residual+ReLU) forward/backward: c=64/256 take the 16-byte kernels, c=4 the generic ones;
    both within bf16 rounding (2e-2) of PyTorch fp32 evaluated on the same bf16-rounded inputs."""
    g = torch.Generator().manual_seed(13)
    n, hw = 4, 6
    x = (torch.randn(n, c, hw, hw, generator=g) * 2 + 0.5).bfloat16().float().requires_grad_(True)
    r = torch.randn(n, c, hw, hw, generator=g).bfloat16().float().requires_grad_(True)
    gamma = (torch.rand(c, generator=g) + 0.5).requires_grad_(True)
    beta = torch.randn(c, generator=g).requires_grad_(True)
    y = F.relu(F.batch_norm(x, None, None, gamma, beta, True, 0.1, 1e-5) + r)
    gy = torch.randn(y.shape, generator=g).bfloat16().float()
    y.backward(gy)
    xd = x.detach().to(DEV).bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    rd = r.detach().to(DEV).bfloat16().contiguous(memory_format=torch.channels_last).requires_grad_(True)
    gd, bd = gamma.detach().to(DEV).requires_grad_(True), beta.detach().to(DEV).requires_grad_(True)
    yd = ops.bn_act(xd, gd, bd, rd, torch.zeros(c, device=DEV), torch.ones(c, device=DEV), True, 0.1, 1e-5, True)
    yd.backward(gy.to(DEV).bfloat16())
    kw = dict(rtol=2e-2, atol=2e-2)
    torch.testing.assert_close(yd.float().cpu(), y.detach(), **kw)
    torch.testing.assert_close(xd.grad.float().cpu(), x.grad, **kw)
    torch.testing.assert_close(rd.grad.float().cpu(), r.grad, **kw)
    torch.testing.assert_close(gd.grad.cpu(), gamma.grad, rtol=2e-2, atol=5e-2)
    torch.testing.assert_close(bd.grad.cpu(), beta.grad, rtol=2e-2, atol=5e-2)


def test_pool_linear_relu(ops):
    g = torch.Generator().manual_seed(8)
    x = torch.randn(5, 64, 28, 28, generator=g).requires_grad_(True)
    y = F.max_pool2d(x, 3, 2, 1)
    gy = torch.randn(y.shape, generator=g)
    y.backward(gy)
    xd = x.detach().to(DEV).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    yd = ops.maxpool3x3s2(xd)
    yd.backward(gy.to(DEV))
    assert torch.equal(yd.cpu(), y.detach())
    torch.testing.assert_close(xd.grad.cpu(), x.grad, rtol=0, atol=1e-6)

    x = torch.randn(7, 512, 2, 2, generator=g).requires_grad_(True)
    w = torch.randn(10, 512, generator=g).requires_grad_(True)
    b = torch.randn(10, generator=g).requires_grad_(True)
    out = F.linear(torch.flatten(F.adaptive_avg_pool2d(x, 1), 1), w, b)
    go = torch.randn(out.shape, generator=g)
    out.backward(go)
    xd = x.detach().to(DEV).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    wd, bd = w.detach().to(DEV).requires_grad_(True), b.detach().to(DEV).requires_grad_(True)
    od = ops.linear(torch.flatten(ops.avgpool(xd), 1), wd, bd)
    od.backward(go.to(DEV))
    kw = dict(rtol=1e-4, atol=1e-5)
    torch.testing.assert_close(od.cpu(), out.detach(), **kw)
    torch.testing.assert_close(xd.grad.cpu(), x.grad, **kw)
    torch.testing.assert_close(wd.grad.cpu(), w.grad, **kw)
    torch.testing.assert_close(bd.grad.cpu(), b.grad, **kw)

    a = torch.randn(4, 64, 7, 7, generator=g).requires_grad_(True)
    c = torch.randn(4, 64, 7, 7, generator=g).requires_grad_(True)
    z = F.relu(F.relu(a) + c)
    gz = torch.randn(z.shape, generator=g)
    z.backward(gz)
    ad = a.detach().to(DEV).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    cd = c.detach().to(DEV).contiguous(memory_format=torch.channels_last).requires_grad_(True)
    zd = ops.add_relu(ops.relu(ad), cd, True)
    zd.backward(gz.to(DEV))
    assert torch.equal(zd.cpu(), z.detach())
    assert torch.equal(ad.grad.cpu(), a.grad) and torch.equal(cd.grad.cpu(), c.grad)


def test_flat_adam_matches_torch_adam_with_clip(ops):
    from iit_b200.nn.optim import FlatAdam
    g = torch.Generator().manual_seed(9)
    shapes = [(64, 3, 7, 7), (64,), (10, 512), (7,)]
    ps = [torch.randn(s, generator=g) for s in shapes]
    ref = [p.clone().requires_grad_(True) for p in ps]
    ropt = torch.optim.Adam(ref, lr=1e-3)
    mine = [torch.nn.Parameter(p.clone().to(DEV)) for p in ps]
    mine[0].data = mine[0].data.contiguous(memory_format=torch.channels_last)
    opt = FlatAdam(mine, lr=1e-3)
    for step in range(4):
        grads = [torch.randn(s, generator=g) * (10.0 if step % 2 == 0 else 0.01) for s in shapes]
        ropt.zero_grad()
        opt.zero_grad()
        for p, gr in zip(ref, grads):
            p.grad = gr.clone()
        for p, gr in zip(mine, grads):
            p.grad.add_(gr.to(DEV))
        torch.nn.utils.clip_grad_norm_(ref, 1.0)
        ropt.step()
        opt.step(max_norm=1.0)
        for p, q in zip(ref, mine):
            torch.testing.assert_close(q.detach().cpu(), p.detach(), rtol=1e-5, atol=1e-7)


def test_tensor_core_path_is_taken_and_matches_cuda_core_path(ops):
    """bf16 3x3 stride-1 convolutions go through the tcgen05/TMA kernels; results must agree with the
    CUDA-core kernels of the same library on identical bf16 operands (both accumulate in fp32; only the
    summation order differs): 2 bf16 ulps of the output scale."""
    from iit_b200.nn.resnet import Conv2d
    g = torch.Generator().manual_seed(11)
    for (c, k, hw, n, ks, stride, pad) in [
            (64, 64, 14, 9, 3, 1, 1), (128, 128, 7, 5, 3, 1, 1), (256, 256, 4, 11, 3, 1, 1), (512, 512, 2, 37, 3, 1, 1),
            (64, 64, 21, 3, 3, 1, 1), (64, 128, 14, 4, 3, 1, 1),
            (64, 128, 14, 9, 3, 2, 1), (128, 256, 7, 5, 3, 2, 1), (256, 512, 4, 11, 3, 2, 1),      # strided 3x3
            (64, 128, 14, 9, 1, 2, 0), (128, 256, 7, 5, 1, 2, 0), (256, 512, 4, 11, 1, 2, 0),      # downsample 1x1
            (64, 128, 21, 2, 3, 2, 1), (64, 128, 21, 2, 1, 2, 0),                                   # odd sizes (84 px)
            # the TMA-store epilogue clips what lies outside the tensor: image counts off the pixel box (1, 129, 257 images
            # against boxes of 8 / 32 / 128 images), grids no power-of-two box tiles (11 x 11, 3 x 3, 6 x 6), token GEMMs
            # whose row count is no multiple of 128, 64 / 192 / 320 filters (one, three, five 64-channel rounds)
            (128, 128, 4, 1, 3, 1, 1), (256, 256, 2, 129, 3, 1, 1), (128, 256, 1, 257, 1, 1, 0), (128, 128, 11, 3, 3, 1, 1),
            (256, 256, 3, 7, 3, 1, 1), (128, 128, 6, 10, 3, 1, 1), (64, 192, 1, 300, 1, 1, 0), (192, 64, 1, 300, 1, 1, 0),
            (64, 320, 1, 77, 1, 1, 0), (64, 64, 14, 1, 3, 1, 1), (64, 64, 14, 150, 3, 1, 1)]:
        conv = Conv2d(c, k, ks, stride, pad).to(DEV)
        ohw = (hw + 2 * pad - ks) // stride + 1
        x = torch.randn(n, c, hw, hw, generator=g).to(DEV).bfloat16().contiguous(memory_format=torch.channels_last)
        gy = torch.randn(n, k, ohw, ohw, generator=g).to(DEV).bfloat16().contiguous(memory_format=torch.channels_last)
        outs = {}
        for use_tc in (True, False):
            ops.USE_TENSOR_CORES = use_tc
            seen = []
            inner = ops.call
            ops.call = lambda name, *a, inner=inner, seen=seen: (seen.append(name), inner(name, *a))[1]
            try:
                xd = x.clone().requires_grad_(True)
                conv.weight.grad = None
                y = conv(xd)
                y.backward(gy)
            finally:
                ops.call = inner
                ops.USE_TENSOR_CORES = True
            assert ("iit_conv2d_fprop_tc" in seen) == use_tc and ("iit_conv2d_dgrad_tc" in seen) == use_tc, seen
            assert ("iit_conv2d_wgrad_tc" in seen) == use_tc, seen
            outs[use_tc] = (y.detach().float(), xd.grad.float(), conv.weight.grad.detach().float().clone())
        for a, b, what in zip(outs[True], outs[False], ("fprop", "dgrad", "wgrad")):
            scale = b.abs().max().item()
            assert (a - b).abs().max().item() <= 2 * 2 ** -8 * scale, (what, c, k, hw, ks, stride)


@pytest.mark.parametrize("hw,layout", [(56, "nchw_f32"), (84, "nchw_f32"), (56, "cl_bf16")])
def test_stem_tensor_core_path(ops, hw, layout):
    """7x7/2 stem through the Conv2d module: forward = space-to-depth + shifted-window tcgen05 convolution, weight
    gradient = shifted-window kernel on the same image (56 px) or im2col + tcgen05 1x1 wgrad with 192 channels (84 px)."""
    from iit_b200.nn.resnet import Conv2d
    g = torch.Generator().manual_seed(12)
    n = 6
    x = torch.rand(n, 3, hw, hw, generator=g).bfloat16().float()
    w = (torch.randn(64, 3, 7, 7, generator=g) / math.sqrt(147)).bfloat16().float().requires_grad_(True)
    y = F.conv2d(x, w, None, 2, 3)
    gy = torch.randn(y.shape, generator=g).bfloat16().float()
    y.backward(gy)
    conv = Conv2d(3, 64, 7, 2, 3).to(DEV)
    with torch.no_grad():
        conv.weight.copy_(w)
    xd = x.to(DEV)
    if layout == "cl_bf16":
        xd = xd.bfloat16().contiguous(memory_format=torch.channels_last)
    seen = []
    inner = ops.call
    ops.call = lambda name, *a: (seen.append(name), inner(name, *a))[1]
    try:
        sums = ops.new_bn_sums(64, DEV)
        yd = conv(xd, sums, torch.bfloat16)
        yd.backward(gy.to(DEV).bfloat16())
    finally:
        ops.call = inner
    assert "iit_stem_s2d" in seen and "iit_stem_conv_s2d" in seen
    if hw <= 58:
        assert "iit_stem_wgrad_s2d" in seen and "iit_stem_im2col" not in seen
    else:
        assert "iit_stem_im2col" in seen and "iit_conv2d_wgrad_tc" in seen
    assert "iit_conv2d_fprop" not in seen and "iit_conv2d_wgrad" not in seen
    for a, b, what in ((yd.float().cpu(), y.detach(), "fprop"), (conv.weight.grad.cpu(), w.grad, "wgrad")):
        scale = b.abs().max().item()
        assert (a - b).abs().max().item() <= 1e-2 * scale, (what, hw, layout)
    rows = yd.numel() // 64
    torch.testing.assert_close(sums[0].cpu().float() / rows, yd.float().cpu().mean(dim=(0, 2, 3)), rtol=1e-3, atol=1e-4)


def test_stem_cols_cache_follows_the_tensor(ops):
    """ops.stem_cols: one im2col per batch tensor; an in-place change of the tensor (new `_version`) rebuilds it into
    the same buffer, a different tensor object gets its own entry, dead tensors are dropped."""
    import gc
    x = torch.rand(3, 3, 56, 56, device=DEV)
    c1 = ops.stem_cols(x)
    assert ops.stem_cols(x) is c1
    assert torch.equal(c1, ops.stem_im2col(x))
    x.mul_(0.5)
    c2 = ops.stem_cols(x)
    assert c2 is c1 and torch.equal(c2, ops.stem_im2col(x))
    y = torch.rand(3, 3, 56, 56, device=DEV)
    assert ops.stem_cols(y) is not c1
    keys = [(id(x), "col"), (id(y), "col")]
    assert all(k in ops._batch_cache for k in keys)
    del x, y
    gc.collect()
    assert not any(k in ops._batch_cache for k in keys)     # (other tests' tensors may come and go: count only these two)


def _conv_desc(n, h, w, c, k, r, stride, pad):
    from iit_b200._lib import Conv
    d = Conv()
    d.n, d.h, d.w, d.c, d.k, d.r, d.s, d.stride, d.pad = n, h, w, c, k, r, r, stride, pad
    d.p = (h + 2 * pad - r) // stride + 1
    d.q = (w + 2 * pad - r) // stride + 1
    return d


@pytest.mark.parametrize("n,h,w", [(5, 14, 14), (3, 21, 21), (2, 9, 13), (150, 14, 14), (1, 40, 40)])
def test_shifted_window_conv_is_bit_identical_to_the_tiled_kernel(ops, n, h, w):
    """layer1's 3x3 64->64 convolutions: the shifted-window tcgen05 kernel (padded image loaded once, taps as row
    offsets; bands when the image is tall) accumulates the same products in the same order as the tiled kernel, so
    fprop, its BatchNorm sums and dgrad must agree bit for bit; both are also checked against PyTorch fp32."""
    import ctypes as C
    from iit_b200 import _lib
    g = torch.Generator().manual_seed(3)
    x = torch.randn(n, h, w, 64, generator=g).bfloat16().to(DEV)
    wt = (torch.randn(64, 3, 3, 64, generator=g) * 0.05).bfloat16().to(DEV)       # KRSC
    wt_crsk = wt.permute(3, 1, 2, 0).contiguous()
    d = _conv_desc(n, h, w, 64, 64, 3, 1, 1)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda t: C.c_void_p(t.data_ptr())
    res = {}
    for on in (1, 0):
        was = _lib.lib.iit_conv_win_enable(on)
        try:
            y = torch.full((n, h, w, 64), 7.0, dtype=torch.bfloat16, device=DEV)
            dx = torch.full((n, h, w, 64), 7.0, dtype=torch.bfloat16, device=DEV)
            sums = torch.zeros(2, 64, dtype=torch.float64, device=DEV)
            _lib.call("iit_conv2d_fprop_tc", p(x), p(wt), p(y), C.byref(d), p(sums[0]), p(sums[1]), st)
            _lib.call("iit_conv2d_dgrad_tc", p(y), p(wt_crsk), p(dx), C.byref(d), st)
            need = int(_lib.lib.iit_conv2d_wgrad_tc_scratch_bytes(C.byref(d)))
            scratch = torch.full((need,), 0x7f, dtype=torch.uint8, device=DEV)
            dw = torch.full((64, 3, 3, 64), 7.0, device=DEV)
            _lib.call("iit_conv2d_wgrad_tc", p(y), p(x), p(dw), C.byref(d), p(scratch), need, st)
            res[on] = (y, dx, sums, dw)
        finally:
            _lib.lib.iit_conv_win_enable(was)
    assert torch.equal(res[1][0], res[0][0])
    assert torch.equal(res[1][1], res[0][1])
    # fp32 partial sums (one ulp of a partial of ~30 is 2e-6) grouped differently and added by atomics in varying order
    torch.testing.assert_close(res[1][2], res[0][2], rtol=1e-5, atol=1e-4)
    ref = F.conv2d(x.float().cpu().permute(0, 3, 1, 2), wt.float().cpu().permute(0, 3, 1, 2), padding=1).permute(0, 2, 3, 1)
    torch.testing.assert_close(res[1][0].float().cpu(), ref, rtol=2e-2, atol=2e-2)
    # weight gradient (dy := y): exact bf16 products, fp32 accumulation in a different order -> 1e-3 of the largest entry
    xr = x.float().cpu().permute(0, 3, 1, 2)
    wr = wt.float().cpu().permute(0, 3, 1, 2).clone().requires_grad_(True)
    F.conv2d(xr, wr, padding=1).backward(res[1][0].float().cpu().permute(0, 3, 1, 2))
    dw_ref = wr.grad.permute(0, 2, 3, 1)
    for on in (1, 0):
        err = (res[on][3].cpu() - dw_ref).abs().max().item()
        assert err <= 1e-3 * dw_ref.abs().max().item(), (on, err, dw_ref.abs().max().item())
    yb = res[1][0].double()
    torch.testing.assert_close(res[1][2][0].cpu(), yb.sum((0, 1, 2)).cpu(), rtol=1e-5, atol=1e-3)
    torch.testing.assert_close(res[1][2][1].cpu(), (yb * yb).sum((0, 1, 2)).cpu(), rtol=1e-5, atol=1e-3)


@pytest.mark.parametrize("n,h,w,c,k,r,stride,pad", [
    (1203, 4, 4, 256, 256, 3, 1, 1),      # layer3 body, ragged image count and an odd number of pixel boxes
    (4750, 2, 2, 512, 512, 3, 1, 1),      # layer4 body: two filter blocks per pixel-box pair
    (1203, 7, 7, 128, 256, 3, 2, 1),      # layer3 entry (stride 2): pair kernel in fprop only
    (1203, 7, 7, 128, 256, 1, 2, 0),      # layer3 downsample 1x1
    (4750, 4, 4, 256, 512, 3, 2, 1),      # layer4 entry: fprop and the parity classes of dgrad (256 produced channels)
])
def test_cta_pair_conv_is_bit_identical_to_the_single_cta_kernel(ops, n, h, w, c, k, r, stride, pad):
    """Layers with a multiple of 256 filters: the cta_group::2 kernel (two SMs per 256 x 256 tile, M = 256 MMAs over both
    CTAs' operands) accumulates the same products in the same k order as the single-CTA kernel, so fprop and dgrad must
    agree bit for bit and the BatchNorm sums to fp32 regrouping; the single-CTA kernel is checked against PyTorch above."""
    import ctypes as C
    from iit_b200 import _lib
    g = torch.Generator().manual_seed(11)
    x = torch.randn(n, h, w, c, generator=g).bfloat16().to(DEV)
    wt = (torch.randn(k, r, r, c, generator=g) * 0.05).bfloat16().to(DEV)
    wt_crsk = wt.permute(3, 1, 2, 0).contiguous()
    d = _conv_desc(n, h, w, c, k, r, stride, pad)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda t: C.c_void_p(t.data_ptr())
    res = {}
    for on in (1, 2, 3, 0):      # automatic tile width, 256-filter tiles, 128-filter tiles, single-CTA kernel
        was = _lib.lib.iit_conv_pair_enable(on)
        try:
            y = torch.full((n, d.p, d.q, k), 7.0, dtype=torch.bfloat16, device=DEV)
            dx = torch.full((n, h, w, c), 7.0, dtype=torch.bfloat16, device=DEV)
            sums = torch.zeros(2, k, dtype=torch.float64, device=DEV)
            _lib.call("iit_conv2d_fprop_tc", p(x), p(wt), p(y), C.byref(d), p(sums[0]), p(sums[1]), st)
            _lib.call("iit_conv2d_dgrad_tc", p(y), p(wt_crsk), p(dx), C.byref(d), st)
            torch.cuda.synchronize()
            res[on] = (y, dx, sums)
        finally:
            _lib.lib.iit_conv_pair_enable(was)
    for on in (1, 2, 3):
        assert torch.equal(res[on][0], res[0][0])
        assert torch.equal(res[on][1], res[0][1])
        torch.testing.assert_close(res[on][2], res[0][2], rtol=1e-5, atol=1e-4)   # fp32 partials, atomics in varying order
    yb = res[1][0].double()
    torch.testing.assert_close(res[1][2][0], yb.sum((0, 1, 2)), rtol=1e-5, atol=1e-3)
    torch.testing.assert_close(res[1][2][1], (yb * yb).sum((0, 1, 2)), rtol=1e-5, atol=1e-3)


@pytest.mark.parametrize("n,hw", [(4, 56), (3, 84), (149, 56), (2, 30)])
def test_stem_space_to_depth_conv_matches_conv2d(ops, n, hw):
    """7x7/2 stem as space-to-depth + 4x4 shifted-window convolution (iit_stem_s2d / iit_stem_filter_s2d /
    iit_stem_conv_s2d) against F.conv2d in fp32 on the bf16-rounded image and filter: products are exact in fp32, so
    the only differences are accumulation order and the bf16 rounding of the result (2^-8 relative)."""
    import ctypes as C
    from iit_b200 import _lib
    g = torch.Generator().manual_seed(5)
    x = torch.rand(n, 3, hw, hw, generator=g)
    w = torch.randn(64, 3, 7, 7, generator=g) * 0.1
    xd, wd = x.to(DEV), w.to(DEV)
    sh = (hw + 6) // 2
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    p = lambda t: C.c_void_p(t.data_ptr())
    s2d = torch.empty(n, sh, sh, 16, dtype=torch.bfloat16, device=DEV)
    w2 = torch.empty(64, 4, 4, 16, dtype=torch.bfloat16, device=DEV)
    _lib.call("iit_stem_s2d", p(xd), _lib.IIT_F32, (C.c_int64 * 4)(*xd.stride()), p(s2d), n, hw, hw, st)
    _lib.call("iit_stem_filter_s2d", p(wd), (C.c_int64 * 4)(*wd.stride()), p(w2), st)
    # the two re-layouts are pure data movement + rounding: exact
    xp = F.pad(x, (3, 3, 3, 3)).bfloat16()
    want = torch.zeros(n, sh, sh, 16, dtype=torch.bfloat16)
    for a in range(2):
        for b in range(2):
            want[..., (a * 2 + b) * 3:(a * 2 + b) * 3 + 3] = xp[:, :, a::2, b::2].permute(0, 2, 3, 1)
    assert torch.equal(s2d.cpu(), want)
    y = torch.full((n, hw // 2, hw // 2, 64), 7.0, dtype=torch.bfloat16, device=DEV)
    sums = torch.zeros(2, 64, dtype=torch.float64, device=DEV)
    _lib.call("iit_stem_conv_s2d", p(s2d), p(w2), p(y), n, sh, sh, p(sums[0]), p(sums[1]), st)
    ref = F.conv2d(x.bfloat16().float(), w.bfloat16().float(), stride=2, padding=3).permute(0, 2, 3, 1)
    torch.testing.assert_close(y.float().cpu(), ref, rtol=1e-2, atol=1e-2)
    yb = y.double()
    torch.testing.assert_close(sums[0].cpu(), yb.sum((0, 1, 2)).cpu(), rtol=1e-5, atol=1e-3)
    torch.testing.assert_close(sums[1].cpu(), (yb * yb).sum((0, 1, 2)).cpu(), rtol=1e-5, atol=1e-3)
    # weight gradient on the same raster (s2d width <= 32, i.e. images up to 58 pixels; wider ones use im2col)
    need = int(_lib.lib.iit_stem_wgrad_s2d_scratch_bytes(n, sh, sh))
    if hw <= 58:
        assert need > 0
        gy = torch.randn(n, hw // 2, hw // 2, 64, generator=g).bfloat16()
        gyd = gy.to(DEV)
        scratch = torch.full((need,), 0x7f, dtype=torch.uint8, device=DEV)
        dw = torch.full((64, 3, 7, 7), 7.0, device=DEV).contiguous(memory_format=torch.channels_last)
        _lib.call("iit_stem_wgrad_s2d", p(gyd), p(s2d), p(dw), (C.c_int64 * 4)(*dw.stride()), n, sh, sh, p(scratch), need, st)
        wr = w.bfloat16().float().requires_grad_(True)
        F.conv2d(x.bfloat16().float(), wr, stride=2, padding=3).backward(gy.float().permute(0, 3, 1, 2))
        err = (dw.cpu() - wr.grad).abs().max().item()
        assert err <= 1e-3 * wr.grad.abs().max().item(), (err, wr.grad.abs().max().item())
    else:
        assert need == 0


def test_backward_adds_parameter_gradients_into_marked_buffers(ops):
    """Parameters whose .grad is the optimizer's flat-buffer view (FlatAdam marks it) get their gradients added in place by
    the backward kernels (conv weight-gradient reduction, iit_bn_param_grads) and autograd receives None for them. Same
    numbers, bit for bit, as the ordinary path where autograd accumulates the returned gradient; a second backward adds."""
    from iit_b200.nn.resnet import BatchNorm2d, Conv2d
    g = torch.Generator().manual_seed(21)
    x = torch.randn(6, 64, 14, 14, generator=g).to(DEV).bfloat16().contiguous(memory_format=torch.channels_last)
    gy = torch.randn(6, 64, 14, 14, generator=g).to(DEV).bfloat16().contiguous(memory_format=torch.channels_last)

    def build(direct):
        torch.manual_seed(5)
        conv, bn = Conv2d(64, 64, 3, 1, 1).to(DEV), BatchNorm2d(64).to(DEV)
        with torch.no_grad():
            bn.weight.uniform_(0.5, 1.5)
            bn.bias.normal_(0, 0.3)
        params = [conv.weight, bn.weight, bn.bias]
        if direct:
            for p in params:
                p.grad = torch.zeros_like(p)
                p._iit_direct_grad = p.grad
                assert ops.direct_grad(p) is p.grad
        return conv, bn, params

    def run(direct, times):
        conv, bn, params = build(direct)
        xr = x.clone().requires_grad_(True)
        for _ in range(times):
            y = bn(conv(xr), relu=True)
            y.backward(gy)
        return [p.grad.clone() for p in params], xr.grad.clone()

    seen = []
    inner = ops.call
    ops.call = lambda name, *a: (seen.append(name), inner(name, *a))[1]
    try:
        gd, dxd = run(True, 1)
    finally:
        ops.call = inner
    assert "iit_conv2d_wgrad_tc_accumulate" in seen and "iit_bn_param_grads" in seen and "iit_conv2d_wgrad_tc" not in seen
    gn, dxn = run(False, 1)
    assert torch.equal(dxd, dxn)
    for a, b in zip(gd, gn):
        assert a.stride() == b.stride() or a.dim() == 1
        assert torch.equal(a, b)
    gd2, _ = run(True, 2)
    gn2, _ = run(False, 2)
    for a, b in zip(gd2, gn2):      # (g + p1) + p2 + ... against g + (p1 + p2 + ...): fp32 rounding order only
        torch.testing.assert_close(a, b, rtol=1e-4, atol=1e-4)


@pytest.mark.parametrize("dtype,c", [(torch.bfloat16, 64), (torch.bfloat16, 256), (torch.float32, 64), (torch.bfloat16, 4)])
def test_bn_relu_bitmask_backward_equals_reading_y(ops, dtype, c):
    """BatchNorm(+residual)+ReLU keeps one bit per output element (y > 0) for its backward instead of having it re-read y:
    same dx, dresidual, dgamma, dbeta bit for bit (c = 4 takes the generic kernels, which ignore the mask)."""
    g = torch.Generator().manual_seed(31)
    x = (torch.randn(5, c, 6, 6, generator=g) * 2).to(DEV).to(dtype).contiguous(memory_format=torch.channels_last)
    r = torch.randn(5, c, 6, 6, generator=g).to(DEV).to(dtype).contiguous(memory_format=torch.channels_last)
    gy = torch.randn(5, c, 6, 6, generator=g).to(DEV).to(dtype).contiguous(memory_format=torch.channels_last)
    gamma, beta = (torch.rand(c, generator=g) + 0.5).to(DEV), torch.randn(c, generator=g).to(DEV)

    def run(use_mask, with_res):
        was = ops.BN_RELU_MASK
        ops.BN_RELU_MASK = use_mask
        try:
            xd, rd = x.clone().requires_grad_(True), r.clone().requires_grad_(True)
            gd, bd = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
            y = ops.bn_act(xd, gd, bd, rd if with_res else None, torch.zeros(c, device=DEV), torch.ones(c, device=DEV), True, 0.1,
                           1e-5, True)
            y.backward(gy)
            return y.detach(), xd.grad, rd.grad, gd.grad, bd.grad
        finally:
            ops.BN_RELU_MASK = was

    for with_res in (False, True):
        a, b = run(True, with_res), run(False, with_res)
        for u, v in zip(a, b):
            assert (u is None) == (v is None)
            if u is not None:
                assert torch.equal(u, v)


def test_batched_filter_relayout_matches_single(ops):
    """iit_filter_relayout_batched (one launch for every filter copy of the model; CRSK copies of channels-last filters go
    through a shared-memory transposition) against iit_filter_relayout entry by entry, ragged sizes included."""
    from iit_b200._lib import IIT_BF16, Relayout
    g = torch.Generator().manual_seed(41)
    filters = [torch.randn(k, c, r, r, generator=g).to(DEV).contiguous(memory_format=torch.channels_last)
               for k, c, r in [(64, 64, 3), (128, 64, 1), (70, 40, 3), (33, 7, 1)]]
    filters.append(torch.randn(48, 24, 3, 3, generator=g).to(DEV))           # NCHW-contiguous: the generic path
    table = (Relayout * (2 * len(filters)))()
    outs = []
    for i, w in enumerate(filters):
        for j in range(2):
            o = torch.full((w.numel(),), 7.0, dtype=torch.bfloat16, device=DEV)
            outs.append(o)
            e = table[2 * i + j]
            e.w, e.out = w.data_ptr(), o.data_ptr()
            for d in range(4):
                e.stride[d] = w.stride(d)
            e.k, e.c, e.r, e.s = w.shape
            e.to_crsk, e.dtype = j, IIT_BF16
    dev_table = torch.frombuffer(bytearray(bytes(table)), dtype=torch.uint8).clone().to(DEV)
    ops.filter_relayout_batched(dev_table, len(outs), max(w.numel() for w in filters))
    for i, w in enumerate(filters):
        for j in range(2):
            ref = ops.filter_relayout(w, bool(j), torch.bfloat16)
            assert torch.equal(outs[2 * i + j], ref), (tuple(w.shape), j)


def test_dgrad_accumulate_and_grad_slots():
    """iit_conv2d_dgrad_tc_accumulate (epilogue stores replaced by 16-byte bf16 reductions) against dgrad + a tensor
    addition, for the three kernel families (shifted-window 64->64, tiled, stride-2 parity classes / 1x1), bit for bit; and
    the whole ResNet's gradients with the residual blocks' GradSlots on and off (no element-wise addition by autograd
    against autograd's own)."""
    import ctypes as C
    from iit_b200 import _lib, ops
    from iit_b200.nn.resnet import ResNet18LL
    g = torch.Generator().manual_seed(3)

    def desc(n, c, hw, k, ks, stride, pad):
        d = _lib.Conv()
        d.n, d.h, d.w, d.c, d.k, d.r, d.s, d.stride, d.pad = n, hw, hw, c, k, ks, ks, stride, pad
        d.p = d.q = (hw + 2 * pad - ks) // stride + 1
        return d

    def ptr(t):
        return C.c_void_p(t.data_ptr())

    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    for (n, c, hw, k, ks, stride, pad) in [(37, 64, 14, 64, 3, 1, 1), (50, 128, 7, 128, 3, 1, 1), (33, 64, 14, 128, 3, 2, 1),
                                            (33, 64, 14, 128, 1, 2, 0), (300, 256, 4, 256, 3, 1, 1), (1200, 256, 4, 256, 3, 1, 1)]:
        d = desc(n, c, hw, k, ks, stride, pad)
        dy = (torch.randn(n, d.p, d.q, k, generator=g) * 0.5).to(DEV).bfloat16()
        w = (torch.randn(c, ks, ks, k, generator=g) / (k * ks * ks) ** 0.5).to(DEV).bfloat16()     # CRSK
        base = torch.randn(n, hw, hw, c, generator=g).to(DEV).bfloat16()
        dx = torch.empty_like(base)
        _lib.call("iit_conv2d_dgrad_tc", ptr(dy), ptr(w), ptr(dx), C.byref(d), st)
        acc = base.clone()
        _lib.call("iit_conv2d_dgrad_tc_accumulate", ptr(dy), ptr(w), ptr(acc), C.byref(d), st)
        assert torch.equal(acc, base + dx), (n, c, hw, k, ks, stride)

    grads = {}
    x = R.synthetic_pvr(64, 5)[0].to(DEV)
    for use in (True, False):
        ops.USE_GRAD_SLOTS = use
        try:
            ll = ResNet18LL(dtype=torch.bfloat16).to(DEV)
            ll.load_state_dict(R.resnet18_state_dict(0))
            # running statistics: nothing in this forward / backward depends on the order of atomic additions, so the two
            # modes must agree BIT FOR BIT on every convolution's weight gradient (batch statistics would add run-to-run
            # rounding noise of their own, which this random-init network amplifies -- tests/test_bf16_rounding_model.py)
            ll.eval()
            seen = []
            inner = ops.call
            ops.call = lambda name, *a: (seen.append(name), inner(name, *a))[1]
            try:
                ll(x).float().square().mean().backward()
            finally:
                ops.call = inner
            grads[use] = {k: p.grad.detach().double().clone() for k, p in ll.named_parameters()}
            assert (seen.count("iit_conv2d_dgrad_tc_accumulate") == 8) == use, seen.count("iit_conv2d_dgrad_tc_accumulate")
        finally:
            ops.USE_GRAD_SLOTS = True
    for k in grads[True]:
        if k.endswith("conv1.mod.weight") or k.endswith("conv2.mod.weight") or k.endswith("downsample.mod.0.mod.weight"):
            assert torch.equal(grads[True][k], grads[False][k]), k
        else:   # BatchNorm / head parameters: fp64 atomic sums, equal to rounding
            torch.testing.assert_close(grads[True][k], grads[False][k], rtol=1e-5, atol=1e-7)


def test_gemm_epilogues_match_separate_kernels():
    """iit_gemm_bias_act_tc / iit_gemm_dgrad_gelu_tc against torch fp32 on the same bf16 operands: bias, residual and
    gelu_new applied to the fp32 accumulator (one rounding), pre-activation output, GELU backward with column sums."""
    import ctypes as C
    from iit_b200 import _lib
    g = torch.Generator().manual_seed(11)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)

    def ptr(t):
        return None if t is None else C.c_void_p(t.data_ptr())

    for rows, n_in, n_out in [(300, 64, 192), (1000, 64, 3072), (777, 3072, 64), (130, 64, 64)]:
        x = torch.randn(rows, n_in, generator=g).to(DEV).bfloat16()
        w = (torch.randn(n_out, n_in, generator=g) / n_in ** 0.5).to(DEV).bfloat16()          # [out][in]
        bias = torch.randn(n_out, generator=g).to(DEV)
        resid = torch.randn(rows, n_out, generator=g).to(DEV).bfloat16()
        lin = x.float() @ w.float().t()
        # bias + residual
        y = torch.empty(rows, n_out, device=DEV, dtype=torch.bfloat16)
        _lib.call("iit_gemm_bias_act_tc", ptr(x), ptr(w), ptr(bias), ptr(resid), ptr(y), None, rows, n_in, n_out, 0, st)
        torch.testing.assert_close(y.float(), lin + bias + resid.float(), rtol=1e-2, atol=2e-2)
        # bias only, no bias
        _lib.call("iit_gemm_bias_act_tc", ptr(x), ptr(w), None, None, ptr(y), None, rows, n_in, n_out, 0, st)
        torch.testing.assert_close(y.float(), lin, rtol=1e-2, atol=2e-2)
        # bias + gelu_new with the pre-activation kept
        pre = torch.empty_like(y)
        _lib.call("iit_gemm_bias_act_tc", ptr(x), ptr(w), ptr(bias), None, ptr(y), ptr(pre), rows, n_in, n_out, 1, st)
        torch.testing.assert_close(pre.float(), lin + bias, rtol=1e-2, atol=2e-2)
        torch.testing.assert_close(y.float(), R.gelu_new(lin + bias), rtol=1e-2, atol=2e-2)
        # GELU backward in the dgrad epilogue: dpre[rows, n_in] = (dy[rows, n_out] @ w_io[n_in, n_out]^T) * gelu'(pre_in)
        dy = (torch.randn(rows, n_out, generator=g) / n_out ** 0.5).to(DEV).bfloat16()
        w_io = w.t().contiguous()                                                            # [in][out]
        pre_in = torch.randn(rows, n_in, generator=g).to(DEV).bfloat16()
        dpre = torch.empty(rows, n_in, device=DEV, dtype=torch.bfloat16)
        sums = torch.zeros(2, n_in, device=DEV, dtype=torch.float64)
        _lib.call("iit_gemm_dgrad_gelu_tc", ptr(dy), ptr(w_io), ptr(pre_in), ptr(dpre), ptr(sums[0]), ptr(sums[1]), rows, n_in, n_out, st)
        p32 = pre_in.float().requires_grad_(True)
        R.gelu_new(p32).backward(dy.float() @ w.float())
        torch.testing.assert_close(dpre.float(), p32.grad, rtol=2e-2, atol=2e-2)
        torch.testing.assert_close(sums[0].float(), dpre.float().sum(0), rtol=1e-3, atol=1e-2)


def test_gelu_backward_with_bias_and_column_sums():
    """iit_gelu_bwd_bias_colsum: dpre = dpost * gelu_new'(h + bias) and the bias gradient in one pass, against torch fp32."""
    from iit_b200 import ops
    g = torch.Generator().manual_seed(4)
    for rows, n in [(37, 128), (1000, 3072), (5, 8)]:
        h = torch.randn(rows, n, generator=g).to(DEV).bfloat16()
        b = torch.randn(n, generator=g).to(DEV)
        dpost = torch.randn(rows, n, generator=g).to(DEV).bfloat16()
        hd, bd = h.clone().requires_grad_(True), b.clone().requires_grad_(True)
        post = ops.bias_gelu_lean(hd, bd, (rows, n))
        post.backward(dpost)
        h32, b32 = h.float().requires_grad_(True), b.clone().requires_grad_(True)
        ref = R.gelu_new(h32 + b32)
        ref.backward(dpost.float())
        torch.testing.assert_close(post.float(), ref.detach(), rtol=1e-2, atol=1e-2)
        torch.testing.assert_close(hd.grad.float(), h32.grad, rtol=2e-2, atol=2e-2)
        torch.testing.assert_close(bd.grad, b32.grad, rtol=2e-2, atol=2e-2 * rows ** 0.5)


def test_transformer_with_gelu_in_the_gemm_epilogues_matches_the_default_path():
    """nn.transformer.FUSE_GELU_IN_GEMM (off by default: slower) keeps working: same logits and gradients to bf16 rounding."""
    from iit_b200.nn import transformer as T
    from iit_b200.tasks.ioi import make_ll_model
    x = R.synthetic_ioi(8, 7, d_vocab=101, n_names=20)[0][0].to(DEV)
    out = {}
    for flag in (False, True):
        T.FUSE_GELU_IN_GEMM = flag
        try:
            ll = make_ll_model(device=DEV, d_mlp=128, d_vocab=101, n_ctx=32, dtype=torch.bfloat16)
            ll.load_state_dict(R.transformer_state_dict(0, d_mlp=128, d_vocab=101, n_ctx=32))
            y = ll(x)
            y.float().square().mean().backward()
            out[flag] = (y.detach().float(), {n: p.grad.float().clone() for n, p in ll.named_parameters() if p.grad is not None})
        finally:
            T.FUSE_GELU_IN_GEMM = False
    assert float((out[True][0] - out[False][0]).norm() / out[False][0].norm()) < 2e-2
    top = max(float(v.norm()) for v in out[False][1].values())
    for n, gref in out[False][1].items():
        if float(gref.norm()) > 1e-3 * top:
            assert float((out[True][1][n] - gref).norm() / gref.norm()) < 5e-2, n


def test_eval_batchnorm_folded_into_the_convolution(ops):
    """iit_conv2d_fprop_tc_affine: conv + evaluation-mode BatchNorm (+ residual) (+ ReLU) in one kernel, on every kernel
    family the ResNet uses, against convolution -> BatchNorm kernel (which rounds the convolution's output to bf16 in between:
    hence 3 bf16 ulps of the output scale); then the whole network in eval mode with and without the fold."""
    from iit_b200.nn import resnet
    from iit_b200.nn.resnet import BatchNorm2d, Conv2d, Wrapped
    g = torch.Generator().manual_seed(23)
    for (c, k, hw, n, ks, stride, pad) in [(64, 64, 14, 150, 3, 1, 1), (128, 128, 7, 300, 3, 1, 1), (256, 256, 4, 1300, 3, 1, 1),
                                           (64, 128, 14, 40, 1, 2, 0), (512, 512, 2, 2100, 3, 1, 1), (128, 256, 7, 600, 3, 2, 1),
                                           (64, 64, 21, 5, 3, 1, 1)]:
        x = torch.randn(n, c, hw, hw, generator=g).to(DEV).bfloat16().contiguous(memory_format=torch.channels_last)
        conv, bn = Wrapped(Conv2d(c, k, ks, stride, pad)).to(DEV), Wrapped(BatchNorm2d(k)).to(DEV)
        with torch.no_grad():
            bn.mod.weight.copy_(torch.rand(k, generator=g) + 0.5)
            bn.mod.bias.copy_(torch.randn(k, generator=g) * 0.3)
            bn.mod.running_mean.copy_(torch.randn(k, generator=g) * 0.2)
            bn.mod.running_var.copy_(torch.rand(k, generator=g) + 0.5)
        bn.mod.eval()
        ohw = (hw + 2 * pad - ks) // stride + 1
        res = torch.randn(n, k, ohw, ohw, generator=g).to(DEV).bfloat16().contiguous(memory_format=torch.channels_last)
        for residual, relu in ((None, False), (None, True), (res, True)):
            out = {}
            for fold in (True, False):
                ops.FOLD_EVAL_BN = fold
                seen = []
                inner = ops.call
                ops.call = lambda name, *a, inner=inner, seen=seen: (seen.append(name), inner(name, *a))[1]
                try:
                    with torch.no_grad():
                        out[fold] = resnet._conv_bn(conv, bn, x, residual=residual, relu=relu).float()
                finally:
                    ops.call = inner
                    ops.FOLD_EVAL_BN = True
                assert ("iit_conv2d_fprop_tc_affine" in seen) == fold and ("iit_bn_act_fwd" in seen) == (not fold), seen
            scale = out[False].abs().max().item()
            assert (out[True] - out[False]).abs().max().item() <= 3 * 2 ** -8 * scale, (c, k, hw, ks, stride, relu, residual is not None)
    # the network: eval-mode logits with and without the fold
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    ll, _, _ = get_alignment("mnist_pvr", {"input_shape": (1, 3, 56, 56)}, device=DEV, dtype=torch.bfloat16)
    ll.load_state_dict(R.resnet18_state_dict(0))
    ll.eval()
    img = R.synthetic_pvr(96, 31)[0].to(DEV)
    logits = {}
    for fold in (True, False):
        ops.FOLD_EVAL_BN = fold
        try:
            with torch.no_grad():
                logits[fold] = ll(img).float()
        finally:
            ops.FOLD_EVAL_BN = True
    rel = float((logits[True] - logits[False]).norm() / logits[False].norm())
    assert rel < 2e-2, rel
    assert float((logits[True].argmax(1) == logits[False].argmax(1)).float().mean()) > 0.97


def test_weight_gradients_added_straight_into_the_optimizer_buffers():
    """iit_conv2d_wgrad_tc_scatter: with FlatAdam's gradient views in place, the transformer's weight-gradient GEMMs add their
    split-K sums into those views (per-head / transposed layouts through the map table) and hand autograd None. Two
    backward passes (so the second ADDS) against the strided-view path: same numbers to fp32 summation order."""
    from iit_b200.nn import transformer as T
    from iit_b200.nn.optim import FlatAdam
    from iit_b200.tasks.ioi import make_ll_model
    from iit_b200 import _lib
    x = R.synthetic_ioi(8, 7, d_vocab=101, n_names=20)[0][0].to(DEV)
    out = {}
    for flag in (False, True):
        T.DIRECT_WEIGHT_GRADS = flag
        try:
            ll = make_ll_model(device=DEV, d_mlp=128, d_vocab=101, n_ctx=32, dtype=torch.bfloat16)
            ll.load_state_dict(R.transformer_state_dict(0, d_mlp=128, d_vocab=101, n_ctx=32))
            opt = FlatAdam(list(ll.parameters()), lr=1e-3)
            opt.zero_grad()
            before = _lib.kernel_variant_counts()
            for rep in range(2):
                y = ll(x)
                (y.float().square().mean() * (1 + rep)).backward()
            torch.cuda.synchronize()
            out[flag] = {n: p.grad.float().clone() for n, p in ll.named_parameters() if p.grad is not None}
        finally:
            T.DIRECT_WEIGHT_GRADS = True
    top = max(float(v.norm()) for v in out[False].values())
    assert top > 0
    checked = 0
    for n, gref in out[False].items():
        if float(gref.norm()) > 1e-4 * top:
            assert float((out[True][n] - gref).norm() / gref.norm()) < 1e-5, n
            checked += 1
    assert checked >= 10
    # the raw entry point, one map that transposes and one per-head map, against iit_conv2d_wgrad_tc
    from iit_b200 import ops
    g = torch.Generator().manual_seed(3)
    Tn, n_in, n_out = 300, 64, 192
    xx = torch.randn(Tn, n_in, generator=g).to(DEV).bfloat16()
    dy = torch.randn(Tn, n_out, generator=g).to(DEV).bfloat16()
    d = _lib.Conv()
    d.n, d.h, d.w, d.c, d.k, d.r, d.s, d.stride, d.pad, d.p, d.q = Tn, 1, 1, n_in, n_out, 1, 1, 1, 0, 1, 1
    import ctypes as C
    need = int(_lib.lib.iit_conv2d_wgrad_tc_scratch_bytes(C.byref(d)))
    scratch = torch.empty(need, dtype=torch.uint8, device=DEV)
    dwb = torch.empty(n_out * n_in, dtype=torch.float32, device=DEV)
    ops.call("iit_conv2d_wgrad_tc", ops._p(dy), ops._p(xx), ops._p(dwb), C.byref(d), ops._p(scratch), need, ops._stream())
    H, dh = 3, 32                                     # W[h][in][dh] <- dwb[(h*dh + e)*n_in + i], over the first 96 outputs
    specs = [((H, n_in, dh), (dh * n_in, 1, n_in), 0), ((n_in, 96), (1, n_in), 96 * n_in)]
    params = [torch.nn.Parameter(torch.zeros(sz, device=DEV)) for sz, _, _ in specs]
    for prm in params:
        prm.grad = torch.full_like(prm, 0.5)
        prm._iit_direct_grad = prm.grad
    cache = {}
    direct = ops.wgrad_direct_maps(specs, params, cache)
    assert direct is not None and direct[0] == "direct" and ops.wgrad_direct_maps(specs, params, cache)[2] is direct[2]
    ops.call("iit_conv2d_wgrad_tc_scatter", ops._p(dy), ops._p(xx), C.byref(d), ops._p(scratch), need, ops._p(direct[2]), direct[3],
             direct[4], ops._stream())
    for prm, (sz, st, off) in zip(params, specs):
        want = torch.as_strided(dwb, sz, st, off) + 0.5
        torch.testing.assert_close(prm.grad, want, rtol=1e-6, atol=1e-6)


def test_fused_hl_and_cross_entropy_kernel():
    """iit_hl_ce_fwd: the HL model's intervened output and the loss against it in one launch, for explicit targets, the
    MNIST-PVR HL and the IOI HL (padded bf16 logits included), against the separate kernels / torch; gradients too."""
    from iit_b200 import ops
    g = torch.Generator().manual_seed(9)
    # explicit targets, fp32 and bf16, narrow and wide rows
    for rows, classes, dtype in [(37, 10, torch.float32), (4096, 10, torch.bfloat16), (9, 50257, torch.float32), (33, 1000, torch.bfloat16)]:
        lg = (torch.randn(rows, classes, generator=g) * 3).to(DEV).to(dtype)
        tgt = torch.randint(0, classes, (rows,), generator=g).to(DEV)
        want = F.cross_entropy(lg.float().requires_grad_(False), tgt)
        l2 = lg.clone().requires_grad_(True)
        got = ops.cross_entropy(l2, tgt)
        assert abs(got.item() - want.item()) <= 1e-5 * abs(want.item()) + 1e-6, (rows, classes, dtype)
        got.backward()
        ref = lg.float().requires_grad_(True)
        F.cross_entropy(ref, tgt).backward()
        torch.testing.assert_close(l2.grad.float(), ref.grad, rtol=2e-2 if dtype == torch.bfloat16 else 1e-4, atol=1e-6)
    # MNIST-PVR HL fused with the loss
    bv, sv = torch.randint(0, 10, (300, 4), generator=g).to(DEV), torch.randint(0, 10, (300, 4), generator=g).to(DEV)
    lg = torch.randn(300, 10, generator=g).to(DEV)
    for var in (-1, 0, 2):
        loss, hl = ops.hl_cross_entropy(lg, ("pvr", bv, sv if var >= 0 else None, var))
        want_hl = ops.hl_pvr(bv, sv if var >= 0 else None, var)
        assert torch.equal(hl, want_hl)
        assert abs(loss.item() - F.cross_entropy(lg, want_hl).item()) < 1e-5
    ops.raise_on_index_error(DEV)
    # IOI HL fused with the loss on padded bf16 logits: columns >= V are ignored and get a zero gradient
    V, Vp = 50257, 50304
    (bx, _, _), names = R.synthetic_ioi(16, 3, one_hot=False)
    (sx, _, _), _ = R.synthetic_ioi(16, 4, one_hot=False)
    bx, sx, names_d = bx.to(DEV), sx.to(DEV), names.to(DEV)
    lg = (torch.randn(16, Vp, generator=g) * 2).to(DEV).bfloat16()
    lg[:, V:] = 50.0                                     # garbage in the padding must not matter
    for node in ("hook_duplicate", "hook_s_inhibition", "hook_name_mover", "all_nodes_hook"):
        l2 = lg.clone().requires_grad_(True)
        loss, am = ops.hl_cross_entropy(l2, ("ioi", bx, sx, node, names_d, V), classes=V)
        want_am = ops.hl_ioi_last_argmax(bx, sx, node, names_d, V)
        assert torch.equal(am, want_am)
        ref = lg[:, :V].float().requires_grad_(True)
        want = F.cross_entropy(ref, want_am)
        assert abs(loss.item() - want.item()) <= 2e-5 * abs(want.item())
        loss.backward()
        want.backward()
        assert float(l2.grad[:, V:].abs().max()) == 0.0
        torch.testing.assert_close(l2.grad[:, :V].float(), ref.grad, rtol=2e-2, atol=1e-7)
    # a target outside [0, classes) sets the error flag (torch raises there)
    bad = torch.tensor([3, 11, 2], device=DEV)
    ops.cross_entropy(torch.randn(3, 10, device=DEV), bad)
    with pytest.raises(IndexError):
        ops.raise_on_index_error(DEV)


def test_patch_inside_the_convolution_matches_the_hook():
    """north_star (a): the interchange on a convolution's hook point applied in that convolution's epilogue
    (iit_conv2d_fprop_tc_patch; BatchNorm sums of the patched values fused; zero gradient inside the box written by the
    BatchNorm backward) against the Python hook calling the standalone patch kernel. Whole-tensor patches skip the convolution.
    Eval mode (no batch statistics): outputs equal bit for bit. Train mode: loss and gradients agree to bf16 rounding, and the
    fused step launches no patch / zeroing / statistics kernel."""
    from iit_b200 import ops
    from iit_b200.model_pairs import StrictIITModelPair
    from iit_b200.model_pairs.nodes import LLNode
    from iit_b200.nn import resnet
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    from iit_b200.utils.index import Ix
    hook = "mod.layer3.mod.1.mod.conv2.hook_point"
    Bn = 64
    base = tuple(t.to(DEV) for t in R.synthetic_pvr(Bn, 100))
    src = tuple(t.to(DEV) for t in R.synthetic_pvr(Bn, 200))
    res = {}
    for fused in (True, False, "again"):      # "again": the hook path a second time -- the run-to-run noise of batch statistics
        resnet.PATCH_IN_LAYER = fused is True
        try:
            out = {}
            for mode in ("q", "c"):
                ll, hl, corr = get_alignment("mnist_pvr", {"mode": mode, "hook_point": hook, "input_shape": (1, 3, 56, 56)}, dtype=torch.bfloat16)
                ll.load_state_dict(R.resnet18_state_dict(0))
                whole = [LLNode("mod.layer2.mod.0.mod.conv2.hook_point", Ix[[None]]), LLNode("mod.layer4.mod.1.mod.conv1.hook_point", Ix[[None]])]
                pair = StrictIITModelPair(hl, ll, corr, training_args={"lr": 1e-3, "nodes_not_in_circuit": whole})
                opt = pair.make_optimizer()
                ll.eval()
                with torch.no_grad():
                    out[mode, "eval"] = [pair.do_intervention(base, src, k)[1].clone() for k in corr] + \
                                        [pair.get_strict_loss_over_batch(base, src, n, pair.loss_fn).clone() for n in whole]
                ll.train()
                seen = []
                inner = ops.call
                ops.call = lambda name, *a: (seen.append(name), inner(name, *a))[1]
                try:
                    node = list(corr.keys())[1]
                    opt.zero_grad()
                    loss = pair.get_IIT_loss_over_batch(base, src, node, pair.loss_fn)
                    loss.backward()
                    g_iit = torch.cat([p.grad.flatten().double() for p in ll.parameters()])
                    opt.zero_grad()
                    sl = pair.get_strict_loss_over_batch(base, src, whole[0], pair.loss_fn)
                    sl.backward()
                    g_strict = torch.cat([p.grad.flatten().double() for p in ll.parameters()])
                finally:
                    ops.call = inner
                out[mode, "train"] = (loss.item(), g_iit, sl.item(), g_strict, seen)
            res[fused] = out
        finally:
            resnet.PATCH_IN_LAYER = True
    for mode in ("q", "c"):
        for a, b in zip(res[True][mode, "eval"], res[False][mode, "eval"]):
            assert torch.equal(a, b), mode
        lf, gf, sf, gsf, seen_f = res[True][mode, "train"]
        lh, gh, sh, gsh, seen_h = res[False][mode, "train"]
        # Train mode: the BatchNorm sums are accumulated with fp64 atomics in an order that differs from run to run; their
        # last-bit differences, re-rounded to bf16, move this random-init network's loss by ~1e-3 and its gradient by
        # several percent between two runs of the SAME path (tests/test_bf16_rounding_model.py). The fused path must sit
        # inside that noise.
        l2, g2, s2, gs2, _ = res["again"][mode, "train"]
        noise_g = max(float((g2 - gh).norm() / gh.norm()), float((gs2 - gsh).norm() / gsh.norm()))
        noise_l = max(abs(l2 - lh) / abs(lh), abs(s2 - sh) / abs(sh))
        print(f"patch in layer, mode {mode}: fused-vs-hook loss {abs(lf - lh) / abs(lh):.2e} / grad {float((gf - gh).norm() / gh.norm()):.2e}; "
              f"hook run-to-run loss {noise_l:.2e} / grad {noise_g:.2e}")
        assert abs(lf - lh) <= max(3 * noise_l, 2e-3) * abs(lh) and abs(sf - sh) <= max(3 * noise_l, 2e-3) * abs(sh), (mode, lf, lh, sf, sh)
        # The floor is the bound DESIGN 5 records for this network's chaotic sensitivity (gradient rel-L2 0.3): the fused
        # path's batch statistics are summed in another order than the separate statistics kernel's (they agree to 2.5e-7,
        # scratch/stress_stats.py), and one run in four that difference -- not a fault: the eval-mode outputs above are
        # bit-identical -- moved the gradient by 0.13 while two runs of the hook path differed by 1e-3.
        assert float((gf - gh).norm() / gh.norm()) < max(3 * noise_g, 0.3), (mode, noise_g)
        assert float((gsf - gsh).norm() / gsh.norm()) < max(3 * noise_g, 0.3), (mode, noise_g)
        assert "iit_conv2d_fprop_tc_patch" in seen_f and "iit_bn_act_bwd_apply_box" in seen_f
        assert not {"iit_patch_fwd", "iit_patch_bwd"} & set(seen_f), set(seen_f) & {"iit_patch_fwd", "iit_patch_bwd"}
        assert "iit_patch_fwd" in seen_h and "iit_conv2d_fprop_tc_patch" not in seen_h
        # the fused IIT term takes no separate statistics pass; the whole-tensor strict patch skips one convolution
        assert seen_f.count("iit_bn_stats") < seen_h.count("iit_bn_stats") or seen_h.count("iit_bn_stats") == 0
        assert seen_f.count("iit_conv2d_fprop_tc") == seen_h.count("iit_conv2d_fprop_tc") - 2


def test_deep_few_tile_gemm_is_split_over_the_reduction():
    """The unembedding's input gradient: dx[1024, 64] = dy[1024, 50304] @ W^T -- 8 output tiles, 786 k-blocks. The tiled
    kernel cuts the reduction into ranges that run as tiles of their own and add their results (bf16 reductions) onto the
    zeroed output; against torch fp32 on the same bf16 operands."""
    import ctypes as C
    from iit_b200 import _lib
    g = torch.Generator().manual_seed(2)
    rows, n_in, n_out = 1024, 64, 50304
    d = _lib.Conv()
    d.n, d.h, d.w, d.c, d.k, d.r, d.s, d.stride, d.pad, d.p, d.q = rows, 1, 1, n_in, n_out, 1, 1, 1, 0, 1, 1
    dy = (torch.randn(rows, n_out, generator=g) * 0.05).to(DEV).bfloat16()
    w_io = (torch.randn(n_in, n_out, generator=g) * 0.1).to(DEV).bfloat16()       # [in][out]
    dx = torch.full((rows, n_in), float("nan"), device=DEV, dtype=torch.bfloat16)
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.call("iit_conv2d_dgrad_tc", C.c_void_p(dy.data_ptr()), C.c_void_p(w_io.data_ptr()), C.c_void_p(dx.data_ptr()), C.byref(d), st)
    want = dy.float() @ w_io.float().t()
    err = float((dx.float() - want).norm() / want.norm())
    assert err < 1e-2, err
    assert float((dx.float() - want).abs().max()) < 3e-2 * float(want.abs().max())


def test_bf16_attention_and_lnpre_vector_kernels_match_the_fp32_kernels():
    """The IOI model's shapes in bf16 (16 tokens, d_head 16, d_model 64) run on kernels with 16-byte loads / stores; same
    arithmetic as the general kernels: compared with those in fp32 on the same (bf16-representable) inputs."""
    from iit_b200 import ops
    g = torch.Generator().manual_seed(6)
    for B in (1, 5, 130):
        qkv = torch.randn(B, 16, 3, 4, 16, generator=g).bfloat16()
        gz = torch.randn(B, 16, 4, 16, generator=g).bfloat16()
        a = qkv.to(DEV).requires_grad_(True)
        za = ops.attention(a)
        za.backward(gz.to(DEV))
        b = qkv.float().to(DEV).requires_grad_(True)
        zb = ops.attention(b)
        zb.backward(gz.float().to(DEV))
        torch.testing.assert_close(za.float(), zb, rtol=1e-2, atol=1e-2)
        torch.testing.assert_close(a.grad.float(), b.grad, rtol=2e-2, atol=2e-2)
    for rows in (1, 33, 16384):
        x = (torch.randn(rows, 64, generator=g) * 2 + 0.5).bfloat16()
        gy = torch.randn(rows, 64, generator=g).bfloat16()
        a = x.to(DEV).requires_grad_(True)
        ya = ops.lnpre(a, 1e-5)
        ya.backward(gy.to(DEV))
        b = x.float().to(DEV).requires_grad_(True)
        yb = ops.lnpre(b, 1e-5)
        yb.backward(gy.float().to(DEV))
        torch.testing.assert_close(ya.float(), yb, rtol=1e-2, atol=1e-2)
        torch.testing.assert_close(a.grad.float(), b.grad, rtol=3e-2, atol=3e-2)
