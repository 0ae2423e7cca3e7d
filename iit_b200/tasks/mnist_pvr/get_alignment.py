"""(ll_model, hl_model, corr) for the PVR tasks (reference: iit/tasks/mnist_pvr/get_alignment.py:9-31,
defaults from iit/tasks/task_loader.py:35-48)."""
from __future__ import annotations

import torch as t

from iit_b200.nn.resnet import ResNet18LL
from iit_b200.tasks.mnist_pvr.pvr_check_leaky_hl import MNIST_PVR_Leaky_HL, get_corr as get_corr_leaky
from iit_b200.tasks.mnist_pvr.pvr_hl import MNIST_PVR_HL, get_corr

DEFAULT_CONFIG = {"mode": "q", "hook_point": "mod.layer3.mod.1.mod.conv2.hook_point", "model": "resnet18",
                  "pad_size": 7}


def get_alignment(task: str, config: dict, device="cuda", dtype=t.float32, ll_model=None):
    cfg = {**DEFAULT_CONFIG, **config}
    if cfg["model"] != "resnet18":
        raise ValueError(f"Unknown model {cfg['model']}")
    if ll_model is None:
        ll_model = ResNet18LL(num_classes=10, dtype=dtype).to(device)
    if task == "mnist_pvr":
        return ll_model, MNIST_PVR_HL(), get_corr(cfg["mode"], cfg["hook_point"], ll_model, cfg["input_shape"])
    if task == "pvr_leaky":
        return ll_model, MNIST_PVR_Leaky_HL(), get_corr_leaky(cfg["mode"], cfg["hook_point"], ll_model, cfg["input_shape"])
    raise ValueError(f"Unknown task {task}")
