"""TEST INFRASTRUCTURE: import-only stub."""


def __getattr__(name):
    raise RuntimeError(f"matplotlib.pyplot.{name}: plotting is unavailable in the oracle environment")
