"""Hook-name helpers for wrapped modules (reference: iit/utils/wrapper.py:69-70)."""


def get_hook_points(model):
    return [k for k in model.hook_dict.keys() if "conv" in k]
