"""CPU tests of the host side: API mirrors (index / nodes / correspondence / node picker / hooks),
the C-ABI library's exported symbols, and a dry run of the model-pair wiring (no kernel executes)."""
import ctypes
import os
import re

import numpy as np
import pytest
import torch

from iit_b200 import _lib
from iit_b200.model_pairs import HLNode, IITBehaviorModelPair, LLNode, StrictIITModelPair
from iit_b200.utils.correspondence import Correspondence
from iit_b200.utils.index import Ix, TorchIndex
from iit_b200.utils import node_picker

from dryrun import dry_run

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_header_symbols_exported():
    """Every function declared in include/iit_b200.h is exported by the .so and bound in _lib."""
    hdr = open(os.path.join(ROOT, "include", "iit_b200.h")).read()
    declared = set(re.findall(r"\b(iit_[a-z0-9_]+)\s*\(", hdr))
    assert declared, "no declarations parsed"
    lib = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(lib, name), f"{name} declared in the header but not exported"
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    assert _lib.lib.iit_version() >= 100


def test_index_reference_cases():  # reference tests/test_index.py
    assert Ix[:, :, :, :] == Ix[:, :, :, :]
    assert Ix[:, :, 0, :] != Ix[:, :, 1, :]
    assert Ix[:, :] != Ix[:, :, 1, :]
    assert hash(Ix[:, :, 0, :]) == hash(Ix[:, :, 0, :])
    assert hash(Ix[:, :, 0, :]) != hash(Ix[:, :, 1, :])
    assert Ix[:, :, :, :].intersects(Ix[:, :, 0, :])
    assert not Ix[:, :, 0, :].intersects(Ix[:, :, 2:3, :])
    assert not Ix[:, :, 2, :].intersects(Ix[:, :, 1, :])
    with pytest.raises(ValueError):
        Ix[:, :, 1].intersects(Ix[:, :, :, :])
    assert repr(Ix[:, :, 1, :]) == "[:, :, 1, :]"
    assert repr(Ix[[None]]) == "[:]"


def test_index_box():
    assert Ix[None, None, :2, 2:4].box((8, 256, 4, 4)) == [(0, 8), (0, 256), (0, 2), (2, 2)]
    assert Ix[:, :, 3, :].box((8, 16, 4, 16)) == [(0, 8), (0, 16), (3, 1), (0, 16)]
    assert Ix[[None]].box((5, 6)) == [(0, 5), (0, 6)]
    assert Ix[:, -1].box((5, 6)) == [(0, 5), (5, 1)]
    assert Ix[:, [0, 2]].box((5, 6)) is None
    t = torch.arange(8 * 6 * 4 * 4).reshape(8, 6, 4, 4)
    idx = Ix[None, 1:3, :2, 2:4]
    spans = idx.box(t.shape)
    ref = t[idx.as_index]
    assert list(ref.shape) == [e for _, e in spans]


def test_nodes_and_correspondence():  # reference tests/test_corr.py
    corr = Correspondence()
    hl, ll = HLNode("hl", -1), LLNode("ll", None)
    corr[hl] = ll
    assert corr[hl] == ll and corr["hl"] == ll
    assert corr.get_suffixes() == {"attn": "attn.hook_result", "mlp": "mlp.hook_post"}
    assert ll.index == Ix[[None]]
    c = Correspondence.make_corr_from_dict({"a": ["blocks.0.attn.hook_z"], "b": ["blocks.1.mlp.hook_post"]},
                                           make_suffixes_from_corr=True)
    assert c.get_suffixes() == {"attn": "attn.hook_z", "mlp": "mlp.hook_post"}
    assert {n.name for n in c["a"]} == {"blocks.0.attn.hook_z"}


def test_node_picker_transformer():  # reference tests/test_node_picker.py:6-29
    class M:
        class cfg:
            n_layers, n_heads = 2, 4
    want = []
    for l in range(2):
        want += [LLNode(f"blocks.{l}.attn.hook_result", Ix[:, :, h, :]) for h in range(4)]
        want.append(LLNode(f"blocks.{l}.mlp.hook_post", Ix[[None]]))
    assert node_picker.get_all_nodes(M) == want


def test_ioi_nodes_not_in_circuit_matches_golden():
    from iit_b200.tasks.ioi import corr
    gold = np.load(os.path.join(ROOT, "tests", "golden", "ioi.npz"))

    class M:
        class cfg:
            n_layers, n_heads = 6, 4
    got = [f"{n.name}|{n.index}" for n in node_picker.get_nodes_not_in_circuit(M, corr)]
    assert got == gold["nodes_not_in_circuit"].tolist()


def test_resnet_names_match_reference_golden():
    from iit_b200.nn.resnet import ResNet18LL
    from iit_b200.utils.wrapper import get_hook_points
    gold = np.load(os.path.join(ROOT, "tests", "golden", "pvr.npz"))
    m = ResNet18LL()
    assert set(m.hook_dict.keys()) == set(gold["hook_points"].tolist())
    assert get_hook_points(m) == gold["conv_hook_points"].tolist()
    assert set(m.state_dict().keys()) == set(gold["state_dict_keys"].tolist())
    from oracle import restate
    missing = m.load_state_dict(restate.resnet18_state_dict(0))
    assert not missing.missing_keys and not missing.unexpected_keys


def test_node_sampling_matches_reference_sequence():
    """Same numpy stream and call order as the reference => same nodes (golden from the reference run)."""
    from iit_b200.tasks.mnist_pvr.pvr_hl import MNIST_PVR_HL, hl_nodes
    gold = np.load(os.path.join(ROOT, "tests", "golden", "pvr.npz"))
    corr = {hl_nodes[h]: {LLNode("x", Ix[[None]])} for h in ["hook_tl", "hook_tr", "hook_bl", "hook_br"]}
    pair = IITBehaviorModelPair(MNIST_PVR_HL(), None, corr)
    assert [str(pair.sample_hl_name()) for _ in range(4)] == gold["q/nodes"].tolist()


def test_dry_run_pvr_train_and_eval_steps():
    from iit_b200.tasks.mnist_pvr.get_alignment import get_alignment
    from oracle import restate
    with dry_run() as calls:
        ll, hl, corr = get_alignment("mnist_pvr", {"input_shape": (1, 3, 56, 56)}, device="cpu")
        assert all(isinstance(k, HLNode) for k in corr)
        pair = StrictIITModelPair(hl, ll, corr, training_args={"lr": 1e-3})
        assert len(pair.nodes_not_in_circuit) == 16
        opt = torch.optim.Adam(ll.parameters(), lr=1e-3)
        base, src = restate.synthetic_pvr(4, 0), restate.synthetic_pvr(4, 1)
        ll.train()
        m = pair.run_train_step(base, src, pair.loss_fn, opt)
        assert set(m) == {"train/iit_loss", "train/behavior_loss", "train/strict_loss"}
        assert all(p.grad is not None and p.grad.shape == p.shape for p in ll.parameters())
        ll.eval()
        with torch.no_grad():
            e = pair.run_eval_step(base, src, pair.loss_fn)
        assert set(e) == {"val/iit_loss", "val/IIA", "val/accuracy"}
        hl_o, ll_o = pair.do_intervention(base, src, "hook_tr")
        assert hl_o.shape == (4,) and ll_o.shape == (4, 10)
    assert "iit_patch_fwd" in calls and "iit_conv2d_wgrad" in calls and "iit_hl_pvr" in calls


def test_device_dataset_pair_rule_and_loader_order_on_the_host():
    """Host logic of DeviceIITDataset (the gather itself needs a GPU): the pair table follows the reference's per-item
    rule (iit_dataset.py:33-35) and the loader walks the items in DataLoader(shuffle=True)'s own order."""
    from iit_b200.utils.iit_dataset import DeviceIITDataset, IITDataset
    x, y, v = torch.arange(40.).view(20, 2), torch.arange(20), torch.arange(80).view(20, 4)
    ds = DeviceIITDataset((x, y, v), (x[:7], y[:7], v[:7]), seed=3)
    rule = []
    for i in range(len(ds)):
        rng = np.random.default_rng(3 * 1000000 + i)
        rule.append([int(rng.choice(20)), int(rng.choice(7))])
    assert ds.pair_table().tolist() == rule
    samples = [(x[i], y[i], v[i]) for i in range(20)]
    ref = IITDataset(samples, samples[:7], seed=3, device="cpu")
    seen = []
    ds.gather_batch = lambda items: seen.append(list(items)) or items
    torch.manual_seed(5)
    order = [i for b in ds.make_loader(6, 0) for i in b]
    torch.manual_seed(5)
    want = []
    for base, abl in ref.make_loader(6, 0):
        want += base[1].tolist()            # y of the base sample == its row number
    assert [rule[i][0] for i in order] == want
    with pytest.raises(RuntimeError):
        DeviceIITDataset((x, y, v), (x, y, v)).gather_batch([0, 1])


def _gather_worker(rank, world, port, q):
    import torch.distributed as dist
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port), RANK=str(rank), WORLD_SIZE=str(world))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from iit_b200.dist import DataParallel
        dp = DataParallel()
        c = torch.full((3, 2, 2), rank + 1, dtype=torch.int64)
        out = dp.allgather_counts(c)
        # gradient averaging before the clip (one all-reduce per step_on_loss) and the IIA count reduction
        lin = torch.nn.Linear(3, 2)
        for p_ in lin.parameters():
            p_.grad = torch.full_like(p_, float(rank + 1))
        dp.allreduce_grads(torch.optim.SGD(lin.parameters(), lr=0.1), lin)
        grads = [p_.grad.flatten().tolist() for p_ in lin.parameters()]
        # two-bucket path: a model that names a tap, an optimizer with one flat gradient buffer; the tail bucket is
        # all-reduced from inside the backward (when the tap's gradient arrives), the head after it
        from iit_b200.hooks import HookedRootModule, HookPoint

        class Two(HookedRootModule):
            def __init__(self):
                super().__init__()
                self.a, self.b, self.mid = torch.nn.Linear(3, 3), torch.nn.Linear(3, 2), HookPoint()
                self.setup()
                self.dp_bucket_tap = ("mid", "b.weight")

            def forward(self, x):
                return self.b(self.mid(self.a(x)))

        class FlatOpt:
            def __init__(self, model):
                ps = list(model.parameters())
                self.flat_g = torch.zeros(sum(p_.numel() for p_ in ps))
                self.offs, o = {}, 0
                for p_ in ps:
                    self.offs[id(p_)] = o
                    p_.grad = self.flat_g[o:o + p_.numel()].view_as(p_)
                    o += p_.numel()

            def offset_of(self, p_):
                return self.offs[id(p_)]

        torch.manual_seed(0)
        two = Two()
        fo = FlatOpt(two)
        dp2 = DataParallel()
        fired = []
        for it in range(2):           # the first call attaches the tap (one message), the second uses both buckets
            fo.flat_g.zero_()
            two(torch.full((4, 3), float(rank + 1))).sum().backward()
            local = fo.flat_g.clone()
            fired.append(getattr(dp2, "_tail", None) is not None)
            dp2.allreduce_grads(fo, two)
            both = [torch.zeros_like(local) for _ in range(world)]
            dist.all_gather(both, local)
            assert torch.allclose(fo.flat_g, torch.stack(both).mean(0)), (it, fo.flat_g, both)
        assert fired == [False, True], fired
        counts = dp.allreduce_counts(torch.tensor([rank + 1, 10], dtype=torch.int64)).tolist()
        shard = dp.shard((torch.arange(10), torch.arange(10) * 2))
        q.put((rank, out.shape, out[:, :, 0].tolist(), grads, counts, [t_.tolist() for t_ in shard]))
    finally:
        dist.destroy_process_group()


def test_data_parallel_plumbing_world2_gloo():
    """World size 2 over gloo: gradients are averaged (before the clip), IIA counts summed, global batches sharded
    evenly, and the sweep's per-batch counts concatenated along the batch axis."""
    import socket
    import torch.multiprocessing as mp
    s = socket.socket(); s.bind(("127.0.0.1", 0)); port = s.getsockname()[1]; s.close()
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    procs = [ctx.Process(target=_gather_worker, args=(r, 2, port, q)) for r in range(2)]
    for p in procs:
        p.start()
    res = sorted(q.get(timeout=120) for _ in procs)
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    for rank, shape, col, grads, counts, shard in res:
        assert tuple(shape) == (3, 4, 2) and col == [[1, 1, 2, 2]] * 3
        assert all(g == [1.5] * len(g) for g in grads)          # mean of 1 and 2 on both ranks
        assert counts == [3, 20]
        assert shard == [list(range(rank * 5, rank * 5 + 5)), [2 * i for i in range(rank * 5, rank * 5 + 5)]]


def test_ctypes_structs_match_the_header(tmp_path):
    """The ctypes mirrors of the POD descriptors (iit_box_t, iit_conv_t, iit_relayout_t) have the size and field offsets the
    C compiler gives the structs of include/iit_b200.h: a field added on one side only would shift every later argument."""
    import ctypes as C
    import shutil
    import subprocess
    from iit_b200 import _lib
    if shutil.which("gcc") is None:
        pytest.skip("no C compiler")
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    src = tmp_path / "sizes.c"
    src.write_text('#include <stdio.h>\n#include <stddef.h>\n#include "iit_b200.h"\nint main(void) {\n'
                   '  printf("%zu %zu %zu\\n", sizeof(iit_box_t), sizeof(iit_conv_t), sizeof(iit_relayout_t));\n'
                   '  printf("%zu %zu %zu %zu\\n", sizeof(iit_hl_spec_t), offsetof(iit_hl_spec_t, base), offsetof(iit_hl_spec_t, n_names),\n'
                   '         offsetof(iit_hl_spec_t, d_vocab));\n'
                   '  printf("%zu %zu %zu %zu\\n", offsetof(iit_relayout_t, stride), offsetof(iit_relayout_t, k),\n'
                   '         offsetof(iit_relayout_t, dtype), offsetof(iit_relayout_t, out_pitch));\n'
                   '  printf("%zu %zu %zu %zu\\n", sizeof(iit_wgrad_map_t), offsetof(iit_wgrad_map_t, size), offsetof(iit_wgrad_map_t, stride),\n'
                   '         offsetof(iit_wgrad_map_t, offset));\n  return 0;\n}\n')
    exe = tmp_path / "sizes"
    subprocess.check_call(["gcc", "-I", os.path.join(root, "include"), str(src), "-o", str(exe)])
    out = subprocess.check_output([str(exe)], text=True).split()
    sizes, hl, offs, wm = [int(v) for v in out[:3]], [int(v) for v in out[3:7]], [int(v) for v in out[7:11]], [int(v) for v in out[11:]]
    assert wm == [C.sizeof(_lib.WgradMap), _lib.WgradMap.size.offset, _lib.WgradMap.stride.offset, _lib.WgradMap.offset.offset]
    assert sizes == [C.sizeof(_lib.Box), C.sizeof(_lib.Conv), C.sizeof(_lib.Relayout)]
    assert hl == [C.sizeof(_lib.HlSpec), _lib.HlSpec.base.offset, _lib.HlSpec.n_names.offset, _lib.HlSpec.d_vocab.offset]
    assert offs == [_lib.Relayout.stride.offset, _lib.Relayout.k.offset, _lib.Relayout.dtype.offset, _lib.Relayout.out_pitch.offset]


def test_direct_grad_only_for_the_marked_flat_view():
    """ops.direct_grad: backward kernels may add into p.grad only while it is the very tensor the flat optimizer installed
    and marked; any other gradient tensor (or none) takes the ordinary autograd path."""
    from iit_b200 import ops
    p = torch.nn.Parameter(torch.zeros(8, 4))
    assert ops.direct_grad(p) is None
    p.grad = torch.zeros(8, 4)
    assert ops.direct_grad(p) is None                      # not marked
    p._iit_direct_grad = p.grad
    assert ops.direct_grad(p) is p.grad
    p.grad = torch.zeros(8, 4)                             # replaced (e.g. by another optimizer): mark is stale
    assert ops.direct_grad(p) is None
    p.grad = None
    assert ops.direct_grad(p) is None


def test_wgrad_direct_maps_table_layout():
    """ops.wgrad_direct_maps: the iit_wgrad_map_t table the scatter reduction reads -- one entry per parameter, shape padded to
    three indices from the left, strides / offset of the [out][in] matrix, destination = the marked gradient view; None as soon
    as one parameter's gradient may not be written in place; rebuilt only when a gradient buffer moved."""
    import ctypes as C
    from iit_b200 import _lib, ops
    H, D, dh, n_in = 4, 64, 16, 64
    specs = [((H, D, dh), (dh * n_in, 1, n_in), 0), ((D, 3072), (1, D), 64 * n_in)]
    params = [torch.nn.Parameter(torch.zeros(sz)) for sz, _, _ in specs]
    cache = {}
    with torch.enable_grad():
        assert ops.wgrad_direct_maps(specs, params, cache) is None          # no marked gradient views yet
        for p in params:
            p.grad = torch.zeros_like(p)
            p._iit_direct_grad = p.grad
        got = ops.wgrad_direct_maps(specs, params, cache)
        assert got is not None and got[0] == "direct" and got[3] == 2 and got[4] == D * 3072
        table = (_lib.WgradMap * 2).from_buffer_copy(bytes(got[2].numpy().tobytes()))
        assert table[0].dst == params[0].grad.data_ptr() and list(table[0].size) == [H, D, dh]
        assert list(table[0].stride) == [dh * n_in, 1, n_in] and table[0].offset == 0
        assert list(table[1].size) == [1, D, 3072] and list(table[1].stride) == [0, 1, D] and table[1].offset == 64 * n_in
        assert ops.wgrad_direct_maps(specs, params, cache)[2] is got[2]      # cached
        params[1].grad = torch.zeros_like(params[1])                         # another tensor, not marked: no in-place writes
        assert ops.wgrad_direct_maps(specs, params, cache) is None
    with torch.no_grad():
        assert ops.wgrad_direct_maps(specs, params, cache) is None
