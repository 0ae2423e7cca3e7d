"""iit_b200: the interchange-intervention training / evaluation step of tkwa/iit on B200 (sm_100a).

Importing the package loads libiit_b200.so (iit_b200/_lib.py) and fails if it is missing:
there is no CPU or eager-PyTorch fallback for the kernels."""
from . import _lib  # noqa: F401  (fail loudly, early)

__version__ = "0.1.0"
