"""TEST INFRASTRUCTURE: import-only stub (iit/utils/eval_ablations.py:10)."""


def export(*a, **k):
    raise RuntimeError("dataframe_image is unavailable in the oracle environment")
