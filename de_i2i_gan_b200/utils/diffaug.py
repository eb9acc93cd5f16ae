"""DiffAugment (defectGAN/utils/diffaug.py:9-75) on the fused CUDA kernels (csrc/augment.cu).

Same entry point and semantics as the reference: ``DiffAugment(x, policy='', channels_first=True)`` with ``policy`` a
comma-separated list of 'color', 'translation', 'cutout', applied in the order given.  The random parameters are drawn with the
SAME torch calls, shapes and order as the reference functions (so a seeded run consumes the RNG identically); the arithmetic -
brightness, saturation, contrast, the zero-padded integer translation and the cutout window - runs in one fused pass per
policy group, with its adjoint as the backward (the op sits between the generator output and the discriminator in the G-step,
defectgan_model.py:200-203).
"""
import torch
from torch.autograd import Function

from .. import _lib, ops

_BITS = {"color": 1, "translation": 2, "cutout": 4}  # dei2i_aug
_ORDER = ("color", "translation", "cutout")           # the fused kernel's own order (= the reference's usual policy string)


def _draw(name, x):
    """The torch RNG calls of diffaug.py:22-58 for one policy entry, on x's device -> dict of per-image parameters."""
    n, _, h, w = x.shape
    dev = x.device
    if name == "color":
        return dict(brightness=torch.rand(n, 1, 1, 1, dtype=x.dtype, device=dev),
                    saturation=torch.rand(n, 1, 1, 1, dtype=x.dtype, device=dev),
                    contrast=torch.rand(n, 1, 1, 1, dtype=x.dtype, device=dev))
    if name == "translation":
        sx, sy = int(h * 0.125 + 0.5), int(w * 0.125 + 0.5)
        return dict(tx=torch.randint(-sx, sx + 1, size=[n, 1, 1], device=dev), ty=torch.randint(-sy, sy + 1, size=[n, 1, 1], device=dev))
    if name == "cutout":
        ch, cw = int(h * 0.5 + 0.5), int(w * 0.5 + 0.5)
        return dict(ox=torch.randint(0, h + (1 - ch % 2), size=[n, 1, 1], device=dev),
                    oy=torch.randint(0, w + (1 - cw % 2), size=[n, 1, 1], device=dev))
    raise KeyError(name)


def pack_params(n, device, drawn):
    """{policy entry: drawn tensors} -> fp32 [N, 8] parameter block of dei2i_diffaug_fwd."""
    p = torch.zeros(n, 8, dtype=torch.float32, device=device)
    p[:, 1] = 1.0
    p[:, 2] = 1.0
    if "color" in drawn:
        d = drawn["color"]
        p[:, 0] = d["brightness"].reshape(n).float() - 0.5
        p[:, 1] = d["saturation"].reshape(n).float() * 2
        p[:, 2] = d["contrast"].reshape(n).float() + 0.5
    if "translation" in drawn:
        p[:, 3] = drawn["translation"]["tx"].reshape(n).float()
        p[:, 4] = drawn["translation"]["ty"].reshape(n).float()
    if "cutout" in drawn:
        p[:, 5] = drawn["cutout"]["ox"].reshape(n).float()
        p[:, 6] = drawn["cutout"]["oy"].reshape(n).float()
    return p


class _DiffAug(Function):
    @staticmethod
    def forward(ctx, x, params, bits):
        ctx.bits = bits
        ctx.save_for_backward(params)
        return ops.diffaug_fwd(x, params, bits)

    @staticmethod
    def backward(ctx, dy):
        (params,) = ctx.saved_tensors
        return ops.diffaug_bwd(dy.contiguous().float(), params, ctx.bits), None, None


def _groups(names):
    """Split the policy list into maximal runs that follow the fused kernel's order color -> translation -> cutout."""
    run, last = [], -1
    for nm in names:
        idx = _ORDER.index(nm)
        if idx <= last:
            yield run
            run = []
        run.append(nm)
        last = idx
    if run:
        yield run


def DiffAugment(x, policy="", channels_first=True, _params=None):
    """diffaug.py:9-19.  ``_params`` (tests): {policy entry: drawn tensors} to use instead of drawing."""
    if not policy:
        return x
    if not channels_first:
        x = x.permute(0, 3, 1, 2)
    names = [p for p in policy.split(",")]
    for nm in names:
        if nm not in _BITS:
            raise KeyError(nm)
    x = x.contiguous().float()
    for run in _groups(names):
        drawn = {nm: (_params[nm] if _params is not None else _draw(nm, x)) for nm in run}
        bits = sum(_BITS[nm] for nm in run)
        x = _DiffAug.apply(x, pack_params(x.shape[0], x.device, drawn), bits)
    if not channels_first:
        x = x.permute(0, 2, 3, 1)
    return x.contiguous()
