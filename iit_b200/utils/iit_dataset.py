"""(base, source) pair sampling and collation (API mirror of the reference's
iit/utils/iit_dataset.py:9-84). Index semantics are kept bit-exact on the host: item `i` draws
base and ablation indices from `default_rng(seed * 1_000_000 + i)` with two `choice` calls.
Collation stacks on the host and moves each field to the device in ONE copy (the reference issues
one `.to(device)` per sample and field, iit_dataset.py:49-51)."""
from __future__ import annotations

import numpy as np
import torch
from torch.utils.data import DataLoader, Dataset


class IITDataset(Dataset):
    def __init__(self, base_data, ablation_data, seed=0, every_combination=False, device=None):
        self.base_data = base_data
        self.ablation_data = ablation_data
        self.seed = seed
        self.every_combination = every_combination
        self.device = device if device is not None else ("cuda" if torch.cuda.is_available() else "cpu")

    def pair_indices(self, index):
        if self.every_combination:
            return index // len(self.ablation_data), index % len(self.ablation_data)
        rng = np.random.default_rng(self.seed * 1000000 + index)
        base_index = rng.choice(len(self.base_data))
        ablation_index = rng.choice(len(self.ablation_data))
        return int(base_index), int(ablation_index)

    def __getitem__(self, index):
        b, a = self.pair_indices(index)
        return self.base_data[b], self.ablation_data[a]

    def __len__(self):
        if self.every_combination:
            return len(self.base_data) * len(self.ablation_data)
        return len(self.base_data)

    @staticmethod
    def get_encoded_input_from_torch_input(xy, device=None):
        x, y, int_vars = zip(*xy)
        fields = [torch.stack([torch.as_tensor(v) for v in f]) for f in (x, y, int_vars)]
        if device is not None:
            fields = [f.pin_memory().to(device, non_blocking=True) if (f.device.type == "cpu" and torch.cuda.is_available()
                                                                       and torch.device(device).type == "cuda")
                      else f.to(device) for f in fields]
        return tuple(fields)

    @staticmethod
    def collate_fn(batch, device=None):
        base_input, ablation_input = zip(*batch)
        return (IITDataset.get_encoded_input_from_torch_input(base_input, device),
                IITDataset.get_encoded_input_from_torch_input(ablation_input, device))

    def make_loader(self, batch_size, num_workers) -> DataLoader:
        return DataLoader(self, batch_size=batch_size, shuffle=True, num_workers=num_workers,
                          collate_fn=lambda x: self.collate_fn(x, self.device))


class DeviceIITDataset(IITDataset):
    """IITDataset over datasets that live on the GPU as stacked tensors: `base_data` / `ablation_data` are
    `(x, y, intermediate_vars)` tuples of device tensors with one row per sample.

    What stays exactly as in the reference: item `i` pairs base/ablation rows drawn from
    `default_rng(seed * 1_000_000 + i)` (two `choice` calls, iit_dataset.py:33-35), and `make_loader` walks the items in
    the order torch's own `DataLoader(shuffle=True)` produces from the global RNG (the loader below IS a DataLoader, over
    item numbers). What changes: the pair table is computed once, and a batch is built by one gather kernel per field
    (ops.gather_rows) instead of per-sample indexing, stacking and `.to(device)` (iit_dataset.py:47-59)."""

    def __init__(self, base_data, ablation_data, seed=0, every_combination=False, device=None):
        super().__init__(base_data, ablation_data, seed, every_combination, device)
        for name, d in (("base_data", base_data), ("ablation_data", ablation_data)):
            if not (isinstance(d, (tuple, list)) and len(d) == 3 and all(isinstance(t, torch.Tensor) for t in d)):
                raise TypeError(f"{name} must be an (x, y, intermediate_vars) tuple of stacked tensors")
        self._n_base, self._n_abl = int(base_data[0].shape[0]), int(ablation_data[0].shape[0])
        self._table = None

    def __len__(self):
        return self._n_base * self._n_abl if self.every_combination else self._n_base

    def pair_indices(self, index):
        if self.every_combination:
            return index // self._n_abl, index % self._n_abl
        rng = np.random.default_rng(self.seed * 1000000 + index)
        return int(rng.choice(self._n_base)), int(rng.choice(self._n_abl))

    def pair_table(self) -> np.ndarray:
        """[len, 2] int64: (base row, ablation row) of every item; computed once (the rule is per item, not per epoch)."""
        if self._table is None:
            n = len(self)
            if self.every_combination:
                i = np.arange(n, dtype=np.int64)
                self._table = np.stack([i // self._n_abl, i % self._n_abl], axis=1)
            else:
                self._table = np.array([self.pair_indices(i) for i in range(n)], dtype=np.int64).reshape(n, 2)
        return self._table

    def __getitem__(self, index):
        b, a = self.pair_indices(index)
        return tuple(t[b] for t in self.base_data), tuple(t[a] for t in self.ablation_data)

    def gather_batch(self, items):
        """(base_input, ablation_input) for a list of item numbers: two index uploads, six gathers."""
        from iit_b200 import ops
        if not self.base_data[0].is_cuda:
            raise RuntimeError("DeviceIITDataset batches are gathered by a CUDA kernel (no CPU fallback exists)")
        rows = self.pair_table()[np.asarray(items, dtype=np.int64)]
        dev = self.base_data[0].device
        idx = torch.from_numpy(np.ascontiguousarray(rows.T)).to(dev, non_blocking=True)   # [2, B]
        base = tuple(ops.gather_rows(t, idx[0]) for t in self.base_data)
        abl = tuple(ops.gather_rows(t, idx[1]) for t in self.ablation_data)
        return base, abl

    def make_loader(self, batch_size, num_workers=0) -> DataLoader:
        return DataLoader(range(len(self)), batch_size=batch_size, shuffle=True, num_workers=0, collate_fn=self.gather_batch)


def train_test_split(dataset, test_size=0.2, random_state=None):
    n = len(dataset)
    split = int(n * test_size)
    gen = None if random_state is None else torch.Generator().manual_seed(random_state)
    return torch.utils.data.random_split(dataset, [n - split, split], generator=gen)
