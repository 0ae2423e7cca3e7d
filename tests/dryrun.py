"""Test helper: run the HOST-side code of iit_b200 on a machine without a GPU by turning every
kernel launch into a no-op (outputs stay uninitialised). Checks wiring only -- shapes, strides,
autograd plumbing, argument marshalling -- never values. Not part of the product."""
import contextlib

import torch


@contextlib.contextmanager
def dry_run():
    from iit_b200 import ops

    saved = (ops.call, ops._require_cuda, ops._stream)
    calls = []

    def fake_call(name, *args):
        calls.append(name)

    ops.call = fake_call
    ops._require_cuda = lambda *a: None
    ops._stream = lambda: None
    try:
        yield calls
    finally:
        ops.call, ops._require_cuda, ops._stream = saved
