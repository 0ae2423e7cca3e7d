"""Dataset that yields every base sample once (API mirror of the reference's iit/utils/eval_datasets.py:7-19)."""
from __future__ import annotations

from iit_b200.utils.iit_dataset import IITDataset


class IITUniqueDataset(IITDataset):
    def __getitem__(self, index):
        return self.base_data[index]

    def __len__(self):
        return len(self.base_data)

    @staticmethod
    def collate_fn(batch, device=None):
        return IITDataset.get_encoded_input_from_torch_input(batch, device)
