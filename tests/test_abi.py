"""CPU-only checks of the drop-in boundary: the shared object loads and exports exactly what include/dei2i.h declares
(no compute calls without a GPU), and the ctypes prototypes cover the whole header."""
import os
import re

import pytest

from conftest import ROOT


def _declared():
    src = open(os.path.join(ROOT, "include", "dei2i.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(dei2i_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as ge
    ge.build()
    from de_i2i_gan_b200 import _lib
    lib = _lib.load()
    names = _declared()
    assert len(names) >= 20
    for n in names:
        assert hasattr(lib, n), f"{n} declared in include/dei2i.h but not exported"
    assert sorted(_lib.SIGNATURES) == names, set(names) ^ set(_lib.SIGNATURES)
    assert lib.dei2i_version() >= 100
    assert lib.dei2i_set_tuning(b"no_such_knob", 1) == -1


def test_compute_fails_loudly_without_gpu():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    from de_i2i_gan_b200 import ops
    with pytest.raises(Exception):
        ops.nchw_to_nhwc(torch.zeros(1, 3, 4, 4))  # CPU tensor: rejected, never silently computed on the host


def test_conv_spec_geometry():
    from de_i2i_gan_b200.ops import ConvSpec
    s = ConvSpec(3, 64, 7, 1, 3, "reflect")
    assert s.cin_ld == 8 and s.kf == 448 and s.out_hw(128, 128) == (128, 128)
    s = ConvSpec(64, 128, 4, 2, 1, "reflect")
    assert s.out_hw(128, 128) == (64, 64) and s.kf == 1024 and s.kd == 2048
    s = ConvSpec(2048, 6, 2)
    assert s.cout_ld == 8 and s.out_hw(2, 2) == (1, 1)
