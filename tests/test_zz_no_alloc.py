"""dei2i_channel_stats_static: the statistics entry point for callers that cannot allocate (its scratch is a file-scope
`static __device__` array inside libdei2i.so, include/dei2i.h).  Called straight through the C-ABI with ctypes and compared with the
existing dei2i_channel_stats on the same input: the two run the same kernels, so the bar is bit-equality (tighter than the 1e-5 the
statistics tests use), on two calls in a row - the second call finds the first call's contents in the shared array.

This file is named to be collected last: a failure here must not hide other tests.  It passes both
alone in a fresh process (first use of the array on the device) and after the rest of the suite."""
import ctypes as C

import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from de_i2i_gan_b200 import _lib
    return _lib.load()


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _reference(lib, x, n, hw, c, per_sample):
    need = lib.dei2i_channel_stats_scratch(n, hw, c, per_sample)
    scratch = torch.empty(need, dtype=torch.float32, device="cuda")
    stats = torch.full(((n if per_sample else 1), 2, c), 7.0, dtype=torch.float32, device="cuda")
    rc = lib.dei2i_channel_stats(_ptr(x), n, hw, c, per_sample, _ptr(stats), _ptr(scratch), need, _stream())
    assert rc == 0, lib.dei2i_last_error().decode()
    return stats, need


@pytest.mark.parametrize("per_sample", [1, 0])
def test_static_scratch_matches_caller_scratch_on_two_calls_in_a_row(per_sample):
    lib = _lib()
    g = torch.Generator(device="cuda").manual_seed(3 + per_sample)
    shapes = [(4, 48 * 40, 64), (3, 17 * 23, 256)]  # the second call is smaller and wider: it reuses a part of what the first left behind
    for n, hw, c in shapes:
        x = (torch.randn(n, hw, c, generator=g, device="cuda") * 1.5 + 0.25).bfloat16().contiguous()
        want, need = _reference(lib, x, n, hw, c, per_sample)
        assert need <= lib.dei2i_channel_stats_static_max_floats()
        got = torch.full_like(want, -3.0)
        before = torch.cuda.memory_allocated()
        rc = lib.dei2i_channel_stats_static(_ptr(x), n, hw, c, per_sample, _ptr(got), _stream())
        assert rc == 0, lib.dei2i_last_error().decode()
        assert torch.cuda.memory_allocated() == before  # nothing went through torch's allocator
        torch.cuda.synchronize()
        assert torch.equal(got, want), float((got - want).abs().max())
        # and against plain torch, at the tolerance of the statistics tests
        xf = x.float()
        dims = (1,) if per_sample else (0, 1)
        mean, var = xf.mean(dim=dims), xf.var(dim=dims, unbiased=False)
        assert torch.allclose(got[:, 0], mean.reshape(got[:, 0].shape), rtol=1e-4, atol=1e-5)
        assert torch.allclose(got[:, 1], var.reshape(got[:, 1].shape), rtol=1e-4, atol=1e-5)


def test_static_scratch_rejects_shapes_beyond_its_maximum():
    lib = _lib()
    n, hw, c = 32, 64 * 64, 2048  # needs ~4.9 M scratch floats, the array holds 1 Mi
    assert lib.dei2i_channel_stats_scratch(n, hw, c, 1) > lib.dei2i_channel_stats_static_max_floats()
    x = torch.zeros(8, dtype=torch.bfloat16, device="cuda")  # never read: the size check comes before any launch
    stats = torch.zeros(n, 2, c, dtype=torch.float32, device="cuda")
    rc = lib.dei2i_channel_stats_static(_ptr(x), n, hw, c, 1, _ptr(stats), _stream())
    assert rc == -2, rc  # DEI2I_ERR_INVALID
    msg = lib.dei2i_last_error().decode()
    assert "DEI2I_STATS_STATIC_MAX_FLOATS" in msg and "dei2i_channel_stats" in msg, msg
    torch.cuda.synchronize()
