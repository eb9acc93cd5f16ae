"""TEST INFRASTRUCTURE ONLY - the CPU oracle with BF16 rounding emulated where the CUDA path stores BF16.

Wraps the leaf ops of oracle/defectgan_ref.py so that every activation and every activation gradient is rounded to
BF16 (weights too), while all arithmetic stays FP32 - the numerical model of the sm_100a path (BF16 storage,
FP32 accumulation).  Used to derive / justify the whole-network gradient tolerances of tests/test_gpu_golden.py:
what the BF16 storage format alone does to a ~30-layer backward chain, independent of any kernel.
"""
import contextlib

import torch

from . import defectgan_ref as R


class _RoundBoth(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        return x.bfloat16().float()

    @staticmethod
    def backward(ctx, g):
        return g.bfloat16().float()


q = _RoundBoth.apply
_OPS = ("conv2d", "batch_norm", "act", "spade", "noise_injection", "conv_weight")


@contextlib.contextmanager
def bf16_storage():
    """Inside: R.<leaf op> outputs (and their gradients) are BF16-rounded."""
    saved = {n: getattr(R, n) for n in _OPS}

    def wrap(f):
        return lambda *a, **k: q(f(*a, **k))
    try:
        for n, f in saved.items():
            setattr(R, n, wrap(f))
        yield
    finally:
        for n, f in saved.items():
            setattr(R, n, f)
