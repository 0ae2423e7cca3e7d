"""IOI task: LL config, correspondence (reference: iit/tasks/ioi/__init__.py:8-37)."""
from iit_b200.utils.correspondence import Correspondence

n_layers = 6
n_heads = 4
d_model = 64
d_head = d_model // n_heads
ioi_cfg = {"n_layers": n_layers, "n_heads": n_heads, "d_model": d_model, "d_head": d_head}
all_attns = [f"blocks.{i}.attn.hook_z" for i in range(n_layers)]
all_mlps = [f"blocks.{i}.mlp.hook_post" for i in range(n_layers)]
corr_dict = {
    "hook_duplicate": [all_attns[0]],
    "hook_s_inhibition": [all_attns[2], all_attns[3]],
    "hook_name_mover": [all_attns[4], all_attns[5]],
    "all_nodes_hook": [*all_mlps[:2]],
}
suffixes = {"attn": "attn.hook_z", "mlp": "mlp.hook_post"}
corr = Correspondence.make_corr_from_dict(corr_dict, suffixes=suffixes, make_suffixes_from_corr=False)

from iit_b200.tasks.ioi.ioi_hl import IOI_HL  # noqa: E402


def make_ll_model(device="cuda", dtype=None, **overrides):
    """The LL transformer of train_ioi.py:30-34 (gpt2 cfg with ioi_cfg applied; no tokenizer needed)."""
    from iit_b200.nn.transformer import HookedTransformerLL
    cfg = {"n_ctx": 1024, "d_mlp": 3072, "d_vocab": 50257, "act_fn": "gelu_new", "normalization_type": "LNPre",
           "init_weights": True, **ioi_cfg, **overrides, "device": str(device)}
    return HookedTransformerLL(cfg, dtype=dtype).to(device)
