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


def train_test_split(dataset, test_size=0.2, random_state=None):
    n = len(dataset)
    split = int(n * test_size)
    gen = None if random_state is None else torch.Generator().manual_seed(random_state)
    return torch.utils.data.random_split(dataset, [n - split, split], generator=gen)
