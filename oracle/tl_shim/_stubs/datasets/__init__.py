"""TEST INFRASTRUCTURE: import-only stub for HuggingFace `datasets` (ioi_dataset_tl.py:14)."""


def load_dataset(*a, **k):
    raise RuntimeError("no network: HuggingFace datasets are unavailable in the oracle environment")
