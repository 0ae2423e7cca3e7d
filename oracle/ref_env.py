"""TEST INFRASTRUCTURE (oracle): make the UNMODIFIED reference importable in this container.

Only usable where /root/reference exists (the build container, not the GPU box). Used by
oracle/make_golden.py and by the CPU tests that pin the restatement against the reference.

What it supplies:
  1. sys.path entries for the transformer_lens shim (oracle/tl_shim) and /root/reference;
  2. seeded synthetic MNIST idx-ubyte files under <scratch>/data/MNIST/raw so that
     `datasets.MNIST("./data", download=True)` (iit/tasks/mnist_pvr/utils.py:5-6) finds files and
     skips the download (torchvision's check is existence-only);
  3. WANDB_MODE=disabled.
"""
from __future__ import annotations

import os
import struct
import sys
import tempfile

import numpy as np

REFERENCE_ROOT = "/root/reference"
_HERE = os.path.dirname(os.path.abspath(__file__))
_SCRATCH = None


def reference_available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_ROOT, "iit"))


def _write_idx_images(path, arr):
    with open(path, "wb") as f:
        f.write(struct.pack(">IIII", 0x00000803, arr.shape[0], arr.shape[1], arr.shape[2]))
        f.write(arr.astype(np.uint8).tobytes())


def _write_idx_labels(path, arr):
    with open(path, "wb") as f:
        f.write(struct.pack(">II", 0x00000801, arr.shape[0]))
        f.write(arr.astype(np.uint8).tobytes())


def synthetic_mnist(n, seed):
    """10 fixed random class templates + per-sample noise."""
    rng = np.random.default_rng(seed)
    templates = np.random.default_rng(1234).integers(0, 256, size=(10, 28, 28))
    labels = rng.integers(0, 10, size=n)
    noise = rng.integers(-32, 33, size=(n, 28, 28))
    imgs = np.clip(templates[labels] + noise, 0, 255).astype(np.uint8)
    return imgs, labels.astype(np.uint8)


def setup(n_train=512, n_test=256):
    """Idempotent. Returns the scratch cwd holding ./data/MNIST/raw."""
    global _SCRATCH
    if not reference_available():
        raise RuntimeError("reference tree not present; ref_env is for the build container only")
    os.environ.setdefault("WANDB_MODE", "disabled")
    shim = os.path.join(_HERE, "tl_shim")
    for p in (REFERENCE_ROOT, shim):
        if p not in sys.path:
            sys.path.insert(0, p)
    # import-only stubs for absent presentation libraries; real installs win (appended last)
    stubs = os.path.join(shim, "_stubs")
    if stubs not in sys.path:
        sys.path.append(stubs)
    if _SCRATCH is None:
        _SCRATCH = tempfile.mkdtemp(prefix="iit_oracle_")
        raw = os.path.join(_SCRATCH, "data", "MNIST", "raw")
        os.makedirs(raw)
        xi, yi = synthetic_mnist(n_train, 0)
        _write_idx_images(os.path.join(raw, "train-images-idx3-ubyte"), xi)
        _write_idx_labels(os.path.join(raw, "train-labels-idx1-ubyte"), yi)
        xi, yi = synthetic_mnist(n_test, 1)
        _write_idx_images(os.path.join(raw, "t10k-images-idx3-ubyte"), xi)
        _write_idx_labels(os.path.join(raw, "t10k-labels-idx1-ubyte"), yi)
    os.chdir(_SCRATCH)
    return _SCRATCH
