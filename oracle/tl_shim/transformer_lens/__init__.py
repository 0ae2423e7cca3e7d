"""TEST INFRASTRUCTURE (oracle): stand-in for the absent third-party `transformer_lens==1.12.1`
so that the unmodified reference under /root/reference can be imported. See hook_points.py."""
from . import hook_points  # noqa: F401
from . import utils  # noqa: F401
from .hook_points import HookPoint, HookedRootModule  # noqa: F401
from .HookedTransformer import HookedTransformer, HookedTransformerConfig  # noqa: F401


class ActivationCache(dict):
    """Name only: the reference uses it as a type annotation (base_model_pair.py:26-27)."""
