"""Low-level model for MNIST-PVR: ResNet-18 (fc -> 10) whose layers are the sm_100a kernels.

Replaces `HookedModuleWrapper(torchvision.models.resnet18())`
(reference: iit/tasks/mnist_pvr/get_alignment.py:11-15, iit/utils/wrapper.py:6-66) and keeps what
callers can see of it:
  * hook-point names `mod.<child>...hook_point` (67 of them, e.g. the default patch site
    `mod.layer3.mod.1.mod.conv2.hook_point`, iit/tasks/task_loader.py:41);
  * state-dict keys `mod.conv1.mod.weight`, `mod.layer1.mod.0.mod.bn1.mod.running_mean`, ...;
  * train-mode BatchNorm semantics (batch statistics, running-stat update with momentum 0.1 and
    unbiased variance, num_batches_tracked).
Internally activations are channels-last ([N,C,H,W] views of NHWC memory) in fp32 or bf16, and
BN + residual + ReLU run as one kernel whenever no hook sits between them; a hook on an inner hook
point (bn / relu) makes that block fall back to the unfused kernels so the hook sees the tensor the
reference would give it.
"""
from __future__ import annotations

import math

import torch
import torch.nn as nn

from .. import ops
from ..hooks import HookedRootModule, HookPoint


class Conv2d(nn.Module):
    def __init__(self, cin, cout, k, stride, pad):
        super().__init__()
        w = torch.empty(cout, cin, k, k)
        nn.init.kaiming_normal_(w, mode="fan_out", nonlinearity="relu")
        self.weight = nn.Parameter(w.contiguous(memory_format=torch.channels_last))  # physical KRSC
        self.stride, self.pad = stride, pad
        self._prep = {}

    def prepared(self, dtype):
        """(KRSC filter in `dtype`, thunk for the CRSK copy used by dgrad); cached per weight version."""
        w = self.weight
        key = (w._version, w.data_ptr(), dtype)
        ent = self._prep.get("ent")
        # under CUDA-graph capture the re-layout must be part of the graph (replays see new weights)
        if ent is None or ent[0] != key or ops.capturing():
            direct = dtype == torch.float32 and w.dim() == 4 and w.permute(0, 2, 3, 1).is_contiguous()
            krsc = w.detach() if direct else ops.filter_relayout(w.detach(), False, dtype)
            ent = [key, krsc, None]
            self._prep["ent"] = ent

        def crsk():
            if ent[2] is None:
                ent[2] = ops.filter_relayout(w.detach(), True, dtype)
            return ent[2]

        if ops.capturing():
            crsk()  # the backward of this forward is captured too: make the copy while weights are current

        return ent[1], crsk

    def is_stem(self):
        return tuple(self.weight.shape[1:]) == (3, 7, 7) and self.stride == 2 and self.pad == 3

    def stem_filter(self, s2d: bool):
        """The stem filter as the kernels want it, cached per weight version: [64,4,4,16] bf16 in the space-to-depth
        grouping (`s2d`), or [K,192] bf16 = the KRSC rows (147 values) zero-padded to three 128-byte rows."""
        w = self.weight
        key = (w._version, w.data_ptr(), s2d)
        ent = self._prep.get("stem")
        if ent is None or ent[0] != key or ops.capturing():
            if s2d:
                wp = ops.stem_filter_s2d(w.detach())
            else:
                wp = torch.zeros((w.shape[0], ops.STEM_K), dtype=torch.bfloat16, device=w.device)
                wp[:, :147] = w.detach().permute(0, 2, 3, 1).reshape(w.shape[0], 147)
            ent = (key, wp)
            self._prep["stem"] = ent
        return ent[1]

    def forward(self, x, bn_sums=None, act_dtype=None):
        if act_dtype == torch.bfloat16 and ops.USE_TENSOR_CORES and self.is_stem():
            s2d = self.weight.shape[0] == 64 and x.shape[2] % 2 == 0 and x.shape[3] % 2 == 0
            return ops.stem_conv(x, self.weight, self.stem_filter(s2d), bn_sums)
        krsc, crsk = self.prepared(x.dtype)
        return ops.conv2d(x, self.weight, krsc, crsk, self.stride, self.pad, bn_sums)


class BatchNorm2d(nn.Module):
    def __init__(self, c, eps=1e-5, momentum=0.1):
        super().__init__()
        self.weight = nn.Parameter(torch.ones(c))
        self.bias = nn.Parameter(torch.zeros(c))
        self.register_buffer("running_mean", torch.zeros(c))
        self.register_buffer("running_var", torch.ones(c))
        self.register_buffer("num_batches_tracked", torch.tensor(0, dtype=torch.long))
        self.eps, self.momentum, self.c = eps, momentum, c

    def forward(self, x, residual=None, relu=False, sums=None):
        if self.training:
            self.num_batches_tracked.add_(1)
        return ops.bn_act(x, self.weight, self.bias, residual, self.running_mean, self.running_var, self.training,
                          self.momentum, self.eps, relu, sums)


class ReLU(nn.Module):
    def forward(self, x):
        return ops.relu(x)


class MaxPool(nn.Module):
    def forward(self, x):
        return ops.maxpool3x3s2(x)


class AvgPool(nn.Module):
    def forward(self, x):
        return ops.avgpool(x)


class Linear(nn.Module):
    def __init__(self, i, o):
        super().__init__()
        self.weight = nn.Parameter(torch.empty(o, i))
        self.bias = nn.Parameter(torch.empty(o))
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        bound = 1 / math.sqrt(i)
        nn.init.uniform_(self.bias, -bound, bound)

    def forward(self, x):
        return ops.linear(x, self.weight, self.bias)


class Wrapped(nn.Module):
    """`mod` + the hook point on its output: gives the reference's `<name>.mod...` / `<name>.hook_point` names."""

    def __init__(self, mod):
        super().__init__()
        self.mod = mod
        self.hook_point = HookPoint()


class _Seq(nn.Module):
    """Numbered children ('0', '1', ...) like nn.Sequential / the reference's wrapped Sequential."""

    def __init__(self, *mods):
        super().__init__()
        for i, m in enumerate(mods):
            self.add_module(str(i), m)

    def __getitem__(self, i):
        return self._modules[str(i)]


class BasicBlock(nn.Module):
    def __init__(self, cin, cout, stride):
        super().__init__()
        self.conv1 = Wrapped(Conv2d(cin, cout, 3, stride, 1))
        self.bn1 = Wrapped(BatchNorm2d(cout))
        self.relu = Wrapped(ReLU())
        self.conv2 = Wrapped(Conv2d(cout, cout, 3, 1, 1))
        self.bn2 = Wrapped(BatchNorm2d(cout))
        self.downsample = None
        if stride != 1 or cin != cout:
            self.downsample = Wrapped(_Seq(Wrapped(Conv2d(cin, cout, 1, stride, 0)), Wrapped(BatchNorm2d(cout))))


def _conv_bn(conv: Wrapped, bn: Wrapped, x, residual=None, relu=False, fuse_stats=True, act_dtype=None):
    """conv -> hook -> bn(+residual)(+relu). BN statistics come from the conv epilogue unless a hook
    can modify the conv output first (the patch site): then they are taken from the hooked tensor."""
    bnm = bn.mod
    sums = None
    if bnm.training and fuse_stats and not conv.hook_point.has_hooks():
        sums = ops.new_bn_sums(bnm.c, x.device)
    y = conv.hook_point(conv.mod(x, sums) if act_dtype is None else conv.mod(x, sums, act_dtype))
    return bnm(y, residual=residual, relu=relu, sums=sums)


class _Body(nn.Module):
    def __init__(self, num_classes):
        super().__init__()
        self.conv1 = Wrapped(Conv2d(3, 64, 7, 2, 3))
        self.bn1 = Wrapped(BatchNorm2d(64))
        self.relu = Wrapped(ReLU())
        self.maxpool = Wrapped(MaxPool())
        cin = 64
        for i, (ch, stride) in enumerate([(64, 1), (128, 2), (256, 2), (512, 2)], start=1):
            blocks = _Seq(Wrapped(BasicBlock(cin, ch, stride)), Wrapped(BasicBlock(ch, ch, 1)))
            setattr(self, f"layer{i}", Wrapped(blocks))
            cin = ch
        self.avgpool = Wrapped(AvgPool())
        self.fc = Wrapped(Linear(512, num_classes))


class ResNet18LL(HookedRootModule):
    """`dtype` is the activation/compute-input type (torch.float32 or torch.bfloat16); parameters,
    BN statistics and accumulation are fp32 either way."""

    def __init__(self, num_classes=10, dtype=torch.float32):
        super().__init__()
        self.mod = _Body(num_classes)
        self.act_dtype = dtype
        self.setup()

    # --- forward ----------------------------------------------------------------------------
    def _block(self, wb: Wrapped, x):
        b: BasicBlock = wb.mod
        # conv1 -> bn1 -> relu
        if b.bn1.hook_point.has_hooks():
            out = b.bn1.hook_point(_conv_bn(b.conv1, b.bn1, x))
            out = b.relu.hook_point(b.relu.mod(out))
        else:
            out = b.relu.hook_point(_conv_bn(b.conv1, b.bn1, x, relu=True))
        # identity path
        identity = x
        if b.downsample is not None:
            ds = b.downsample
            identity = ds.mod[1].hook_point(_conv_bn(ds.mod[0], ds.mod[1], x))
            identity = ds.hook_point(identity)
        # conv2 -> bn2 -> (+identity) -> relu
        if b.bn2.hook_point.has_hooks():
            out = b.bn2.hook_point(_conv_bn(b.conv2, b.bn2, out))
            out = ops.add_relu(out, identity, True)
        else:
            out = _conv_bn(b.conv2, b.bn2, out, residual=identity, relu=True)
        out = b.relu.hook_point(out)
        return wb.hook_point(out)

    def refresh_input_cache(self, x):
        """Rebuild the per-batch stem operand of a registered (static) input tensor after its contents changed;
        model pairs call this before replaying captured graphs (ops.stem_cols)."""
        ops.refresh_batch_operands(x)

    def forward(self, x):
        m = self.mod
        stem_tc = self.act_dtype == torch.bfloat16 and ops.USE_TENSOR_CORES and m.conv1.mod.is_stem()
        if not stem_tc and (x.dtype != self.act_dtype or not ops._is_cl(x)):
            # the tensor-core stem reads the caller's tensor (fp32 NCHW included) directly; otherwise convert
            x = x.to(self.act_dtype).contiguous(memory_format=torch.channels_last)
        bn = m.bn1.mod
        fuse_tail = (not m.bn1.hook_point.has_hooks() and not m.relu.hook_point.has_hooks()
                     and bn.c % 64 == 0 and (bn.training or not torch.is_grad_enabled()))
        if fuse_tail:
            # conv1 -> hook -> [bn1 -> relu -> maxpool] in one kernel: nothing between them is observed
            sums = None
            if bn.training and not m.conv1.hook_point.has_hooks():
                sums = ops.new_bn_sums(bn.c, x.device)
            y0 = m.conv1.hook_point(m.conv1.mod(x, sums, self.act_dtype))
            if bn.training:
                bn.num_batches_tracked.add_(1)
            x = ops.bn_relu_maxpool(y0, bn.weight, bn.bias, bn.running_mean, bn.running_var, bn.training, bn.momentum,
                                    bn.eps, sums)
            x = m.maxpool.hook_point(x)
        else:
            if m.bn1.hook_point.has_hooks():
                x = m.bn1.hook_point(_conv_bn(m.conv1, m.bn1, x, act_dtype=self.act_dtype))
                x = m.relu.hook_point(m.relu.mod(x))
            else:
                x = m.relu.hook_point(_conv_bn(m.conv1, m.bn1, x, relu=True, act_dtype=self.act_dtype))
            x = m.maxpool.hook_point(m.maxpool.mod(x))
        for i in range(1, 5):
            layer: Wrapped = getattr(m, f"layer{i}")
            x = self._block(layer.mod[0], x)
            x = self._block(layer.mod[1], x)
            x = layer.hook_point(x)
        x = m.avgpool.hook_point(m.avgpool.mod(x))
        x = torch.flatten(x, 1)
        return m.fc.hook_point(m.fc.mod(x))
