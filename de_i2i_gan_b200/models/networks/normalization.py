"""SPADE (models/networks/normalization.py:10-37) on the sm_100a kernels.  The module tree (param_free_norm,
mlp_shared.0, mlp_gamma, mlp_beta) and therefore the state-dict keys are the reference's; nn.Conv2d objects are
parameter containers only - forward runs the fused path:
    seg --3x3 conv + bias + ReLU (nearest resize folded into the gather)--> actv
    actv --ONE 3x3 conv with N = 2C (gamma | beta)--> gb
    (x, gb) --instance-norm statistics + modulation (+ ReLU, + 2x nearest up-sampling of x)--> out
"""
import os

import torch
from torch import nn

from ... import functional as Fn
from ...ops import ConvSpec


# Exact evaluation of gamma|beta on the 5x5 border-class map when the label map is 1x1 (False: always the dense convs)
CONSTANT_LABEL_CLASSES = os.environ.get("DEI2I_SPADE_DENSE", "0") != "1"


class SPADE(nn.Module):
    def __init__(self, label_nc, norm_nc, hidden_nc=128, kernel_size=(3, 3), padding="same", norm_layer=nn.BatchNorm2d):
        super().__init__()
        self.param_free_norm = norm_layer(norm_nc, affine=False)
        if not isinstance(self.param_free_norm, nn.InstanceNorm2d):
            raise NotImplementedError("the CUDA path implements SPADE over InstanceNorm2d (the DefectGAN configuration)")
        self.mlp_shared = nn.Sequential(nn.Conv2d(label_nc, hidden_nc, kernel_size=kernel_size, padding=padding),
                                        nn.ReLU(inplace=True))
        self.mlp_gamma = nn.Conv2d(hidden_nc, norm_nc, kernel_size=kernel_size, padding=padding)
        self.mlp_beta = nn.Conv2d(hidden_nc, norm_nc, kernel_size=kernel_size, padding=padding)
        k = kernel_size[0] if isinstance(kernel_size, (tuple, list)) else kernel_size
        pad = (k - 1) // 2 if padding == "same" else int(padding)
        self._shared_spec = ConvSpec(label_nc, hidden_nc, k, 1, pad, "zeros")
        self._gb_spec = ConvSpec(hidden_nc, 2 * norm_nc, k, 1, pad, "zeros")

    def forward(self, x, segmap, up=1, relu=False, out_pad=0):
        """x: NCHW tensor (drop-in) or Nhwc; segmap: [B,label_nc,h,w] NCHW or Nhwc.  ``up`` / ``relu`` are the fused
        followers used by NormConvBlock / NormResBlock (architecture.py:242-244,347-348); defaults = reference SPADE.
        ``out_pad`` > 0 asks for the reflect padding of the consuming convolution to be written with the output; the returned
        carrier's ``.pad`` says how much of it was (0 on the paths that do not fuse it)."""
        plain = not isinstance(x, Fn.Nhwc)
        xa = Fn.to_nhwc(x)
        seg = Fn.to_nhwc(segmap)
        n, _, hx, wx = xa.shape
        h, w = hx * up, wx * up
        sconv = self.mlp_shared[0]
        if CONSTANT_LABEL_CLASSES and seg.t.shape[1] == 1 and seg.t.shape[2] == 1 and h >= 5 and w >= 5 \
                and self._shared_spec.k == 3 and self._shared_spec.pad == 1:
            # 1x1 label map (every training call, defectgan_model.py:185-205): the nearest-resized map is constant, so
            # mlp_shared -> mlp_gamma|mlp_beta are evaluated on the 5x5 border-class map (exact; dei2i_spade_cls_fwd)
            if seg.cls5 is None:
                seg.cls5 = Fn.Nhwc(seg.t.expand(n, 5, 5, seg.t.shape[3]).contiguous(), seg.c)
            # gamma|beta depend on the label tensor and this module's weights only: the four generator passes of a G-step use
            # two label tensors (df_labels, nm_labels, nm_labels, df_labels - defectgan_model.py:185-205), so inside one
            # weight_cache() scope (= one loss computation, weights fixed) each is evaluated once and shared; autograd sums
            # the gradients of the passes that used it before the one backward through the two small convolutions
            ck = None
            if Fn._cache is not None and seg.key is not None:
                ck = ("gb5", id(self), seg.key, n, self.training, torch.is_grad_enabled())
            gb5 = Fn._cache.get(ck) if ck is not None else None
            if gb5 is None:
                actv = Fn.conv(seg.cls5, sconv, self._shared_spec, self.training, bias=sconv.bias, act="relu")
                gb5 = Fn.conv_cat(actv, self.mlp_gamma, self.mlp_beta, self._gb_spec)
                if ck is not None:
                    Fn._cache[ck] = gb5
            pad = 0 if plain else out_pad
            out = Fn.spade_modulate_cls(xa, gb5, up=up, relu=relu, eps=self.param_free_norm.eps, out_pad=pad)
            out.pad = pad
            return Fn.from_nhwc(out) if plain else out
        actv = Fn.conv(seg, sconv, self._shared_spec, self.training, bias=sconv.bias, act="relu", logical_hw=(h, w))
        gb = Fn.conv_cat(actv, self.mlp_gamma, self.mlp_beta, self._gb_spec)
        out = Fn.spade_modulate(xa, gb, up=up, relu=relu, eps=self.param_free_norm.eps)
        return Fn.from_nhwc(out) if plain else out
