"""TEST INFRASTRUCTURE (oracle): the few `transformer_lens.utils` helpers the reference touches."""


def get_act_name(name, layer=None, layer_type=None):
    if layer is None:
        return name
    if layer_type:
        return f"blocks.{layer}.{layer_type}.hook_{name}"
    return f"blocks.{layer}.hook_{name}"
