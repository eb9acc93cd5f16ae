"""Building blocks of the DefectGAN networks on the sm_100a kernels.

Mirror of /root/reference/defectGAN/models/networks/architecture.py: same class names, constructor signatures,
sub-module names and registration order (=> identical ``state_dict`` keys, ``init_weights`` RNG order and optimizer
parameter order).  ``nn.Conv2d`` / ``nn.BatchNorm2d`` / ``spectral_norm`` objects are kept as parameter containers;
every ``forward`` body runs the fused CUDA path through ``de_i2i_gan_b200.functional``.

Every block accepts either a plain NCHW tensor (drop-in use: converted at the block edge, result returned as NCHW
FP32) or an ``Nhwc`` carrier (how the networks chain blocks without layout round trips).
"""
import logging
import os

import torch
from torch import nn
from torch.nn.utils import spectral_norm  # old hook API, as the reference (architecture.py:7)

from ... import functional as Fn
from ...ops import ConvSpec
from .normalization import SEAN, SPADE, AdaIN

FUSE_REFLECT_PAD = os.environ.get("DEI2I_FUSE_REFLECT_PAD", "1") != "0"
ROWPACK_STEM = os.environ.get("DEI2I_ROWPACK_STEM", "1") != "0"
_ACT_NAMES = {nn.LeakyReLU: "leaky_relu", nn.ReLU: "relu", nn.Sigmoid: "sigmoid", nn.Tanh: "tanh", nn.Identity: None}


def get_act_layer(act_str):
    """architecture.py:12-26."""
    if act_str == "leaky_relu":
        return nn.LeakyReLU(0.2, inplace=True)
    if act_str == "relu":
        return nn.ReLU(inplace=True)
    if act_str == "sigmoid":
        return nn.Sigmoid()
    if act_str == "tanh":
        return nn.Tanh()
    if act_str is None:
        logging.info("create conv block without activation layer")
        return nn.Identity()
    raise NameError(f"activation layer named {act_str} not defined")


def _first(v):
    return v[0] if isinstance(v, (tuple, list)) else v


def _spec_of(conv: nn.Conv2d) -> ConvSpec:
    """ConvSpec from an nn.Conv2d container (kernel, stride, padding incl. 'same'/'valid', padding_mode)."""
    kh, kw = conv.kernel_size
    pad = conv.padding
    if isinstance(pad, str):
        if pad == "same":
            if kh % 2 == 0 or kw % 2 == 0:
                # asymmetric 'same' padding of even kernels (WGanGenerator): the zero frame is materialised by
                # Fn.upsample_pad (see same_frame), the convolution itself is 'valid'
                if conv.padding_mode != "zeros" or kh != kw:
                    raise NotImplementedError("'same' padding of even kernels: square filters with zero padding only")
                pad = 0
            else:
                pad = (kh - 1) // 2
        else:
            pad = 0
    else:
        pad = _first(pad)
    if conv.padding_mode not in ("zeros", "reflect"):
        raise NotImplementedError(f"padding_mode {conv.padding_mode}")
    if conv.groups != 1 or _first(conv.dilation) != 1:
        raise NotImplementedError("grouped / dilated convolutions")
    return ConvSpec(conv.in_channels, conv.out_channels, kh, _first(conv.stride), pad, conv.padding_mode, kw if kw != kh else 0)


def same_frame(conv: nn.Conv2d):
    """(top, left, bottom, right) zero frame of nn.Conv2d(padding='same') for even kernels - torch pads total = k - 1,
    total // 2 before and the remainder after - else None (symmetric padding is part of the ConvSpec)."""
    kh, kw = conv.kernel_size
    if conv.padding == "same" and kh % 2 == 0:
        total = kh - 1
        lo, hi = total // 2, total - total // 2
        return (lo, lo, hi, hi)
    return None


# ---------------------------------------------------------------------------------------------- noise
_noise_source = None


def set_noise_source(fn):
    """Test hook: ``fn(n, h, w, device) -> fp32 [n,1,h,w]`` replaces the torch RNG draw (None restores it)."""
    global _noise_source
    _noise_source = fn


def draw_noise(n, h, w, device):
    """Same call, shape and order as the reference (architecture.py:387-388): one N(0,1) map per sample."""
    if _noise_source is not None:
        return _noise_source(n, h, w, device).to(device=device, dtype=torch.float32).contiguous()
    return torch.empty(n, 1, h, w, device=device, dtype=torch.float32).normal_()


class NoiseInjection(nn.Module):
    """architecture.py:374-389.  Inside the blocks the addition is fused into the conv epilogue; a stand-alone call on a
    plain tensor is the literal formula."""

    def __init__(self, weight_type="constant", nc=None):
        super().__init__()
        if weight_type == "constant":
            self.weight = nn.Parameter(torch.zeros(1, 1, 1, 1))
        elif weight_type == "vector":
            raise NotImplementedError("per-channel noise weights are not used by DefectGAN and not on the CUDA path")
        else:
            raise NameError(f"weight type named {weight_type} not defined")

    def forward(self, image, noise=None):
        if noise is None:
            b, _, h, w = image.shape
            noise = draw_noise(b, h, w, image.device)
        return image + self.weight * noise


def _noise_args(noise_mod, n, h, w, device):
    if isinstance(noise_mod, NoiseInjection):
        return draw_noise(n, h, w, device), noise_mod.weight
    return None, None


# ---------------------------------------------------------------------------------------------- blocks
def _apply_sn(blocks, use_spectral):
    if use_spectral:
        for i, layer in enumerate(blocks):
            if isinstance(layer, (nn.Conv2d, nn.Linear)):
                blocks[i] = spectral_norm(layer)


class ConvBlock(nn.Module):
    """architecture.py:79-118: [spectral] Conv2d -> [BatchNorm2d | InstanceNorm2d] -> activation."""

    def __init__(self, f_in, f_out, kernel_size=(3, 3), stride=(1, 1), padding=0, padding_mode="zeros", bias=False,
                 norm_layer=None, act_layer=None, use_spectral=False):
        super().__init__()
        blocks = [nn.Conv2d(f_in, f_out, kernel_size=kernel_size, stride=stride, padding=padding,
                            padding_mode=padding_mode, bias=bias)]
        if norm_layer is not None:
            blocks.append(norm_layer(f_out))
        blocks.append(get_act_layer(act_layer))
        _apply_sn(blocks, use_spectral)
        self.conv_block = nn.Sequential(*blocks)
        self._spec = _spec_of(self.conv_block[0])
        self._act = act_layer
        self._has_norm = norm_layer is not None
        self._inorm = self._has_norm and isinstance(self.conv_block[1], nn.InstanceNorm2d)
        if self._has_norm and not isinstance(self.conv_block[1], (nn.BatchNorm2d, nn.InstanceNorm2d)):
            raise NotImplementedError("ConvBlock on the CUDA path supports norm_layer in (None, nn.BatchNorm2d, nn.InstanceNorm2d)")
        if self._inorm:
            inorm = self.conv_block[1]
            if inorm.track_running_stats or act_layer not in _IN_SLOPE:
                raise NotImplementedError("ConvBlock + InstanceNorm2d on the CUDA path: per-image statistics, (leaky) ReLU or no activation")

    @property
    def _norm(self):  # not an attribute: registering the sub-module twice would duplicate its state-dict keys
        return self.conv_block[1] if self._has_norm else None

    def forward(self, x, seg=None, residual=None, carrier_out=False, out_pad=0):
        """out_pad (carrier output with a BatchNorm only): the reflect padding of the NEXT convolution, written by the
        BatchNorm kernel itself (`reflect_pad_of(next_block)`); the result then has ``.pad == out_pad``."""
        plain = not isinstance(x, Fn.Nhwc)
        conv = self.conv_block[0]
        if self._inorm:
            # StyleExtractor (extractor.py:60-69): conv -> InstanceNorm2d -> activation [+ residual]; normalisation, affine (if any)
            # and activation are ONE pass of the AdaIN kernel with per-image scale / shift rows
            xa = Fn.to_nhwc(x)
            y = Fn.conv(xa, conv, _consumer_spec(self._spec, xa), conv.training, bias=conv.bias)
            inorm, n = self.conv_block[1], xa.shape[0]
            one = inorm.weight if inorm.affine else torch.ones(self._spec.cout, device=y.t.device)
            zero = inorm.bias if inorm.affine else torch.zeros(self._spec.cout, device=y.t.device)
            y = Fn.adain_act(y, one.float().expand(n, -1), zero.float().expand(n, -1), slope=_IN_SLOPE[self._act], eps=inorm.eps)
            if residual is not None:
                y = Fn.add_scaled(y, residual, 1.0)
            return y if (carrier_out or not plain) else _edge(y)
        if plain and ROWPACK_STEM and self._norm is not None and conv.bias is None and Fn.rowpack_applicable(self._spec):
            # generator stem (3->64, 7x7 reflect): row-packed image -> 7x1 convolution over 64 channels on the TMA kernels
            y = Fn.rowpack_conv(x, conv, self._spec, conv.training, want_stats=_bn_stats(self._norm))
            y = Fn.batch_norm(y, self._norm, act=self._act, residual=residual, out_pad=out_pad if carrier_out else 0)
            return y if carrier_out else _edge(y)
        xa = Fn.to_nhwc(x)
        if carrier_out:
            plain = False
        if plain:
            out_pad = 0
        spec = _consumer_spec(self._spec, xa)
        if self._norm is None:
            assert residual is None
            # no norm layer: the convolution epilogue itself writes the next convolution's reflect padding (granted on the TMA kernel)
            y = Fn.conv(xa, conv, spec, conv.training, bias=conv.bias, act=self._act,
                        out_fp32=self._act in ("tanh", "sigmoid"), out_pad=out_pad)
            if y.t.dtype != Fn.ops.ACT_DTYPE and not plain:
                y = Fn.Nhwc(y.t.to(Fn.ops.ACT_DTYPE), y.c)
        else:
            y = Fn.conv(xa, conv, spec, conv.training, bias=conv.bias, want_stats=_bn_stats(self._norm))
            y = Fn.batch_norm(y, self._norm, act=self._act, residual=residual, out_pad=out_pad)
        return _edge(y) if plain else y

    def reflect_pad_in(self) -> int:
        """Reflect padding this block's convolution wants on its input (what a producer may pre-write), else 0."""
        return _fusable_pad(self._spec)


_IN_SLOPE = {None: 1.0, "relu": 0.0, "leaky_relu": 0.2}  # the AdaIN kernel's activation is LeakyReLU(slope)


def _bn_stats(bn) -> str:
    """'batch' when the BatchNorm that follows a convolution normalises with batch statistics (the epilogue then emits them)."""
    return "batch" if (bn.training or not bn.track_running_stats) else None


def _edge(y: Fn.Nhwc) -> torch.Tensor:
    if y.t.dtype == torch.float32:
        return y.t[..., : y.c].permute(0, 3, 1, 2).contiguous()
    return Fn.from_nhwc(y)


class DeConvBlock(nn.Module):
    """architecture.py:29-76: [Upsample] -> Conv2d -> [NoiseInjection] -> [norm] -> activation.  DefectGAN uses it for
    the two generator heads only (no up-sampling, noise or norm: generator.py:224-241)."""

    def __init__(self, f_in, f_out, kernel_size=(3, 3), stride=(1, 1), padding=0, padding_mode="zeros", bias=False,
                 up_scale=True, norm_layer=None, act_layer=None, use_spectral=False, add_noise=False):
        super().__init__()
        blocks = []
        if up_scale:
            blocks.append(nn.Upsample(scale_factor=2))
        blocks.append(nn.Conv2d(f_in, f_out, kernel_size=kernel_size, stride=stride, padding=padding,
                                padding_mode=padding_mode, bias=bias))
        if add_noise:
            blocks.append(NoiseInjection())
        if norm_layer is not None:
            blocks.append(norm_layer(f_out))
        blocks.append(get_act_layer(act_layer))
        _apply_sn(blocks, use_spectral)
        self.de_conv_block = nn.Sequential(*blocks)
        self._act = act_layer
        self._conv_idx = next(i for i, m in enumerate(self.de_conv_block) if isinstance(m, nn.Conv2d))
        self._simple = not (up_scale or add_noise or norm_layer is not None)
        self._spec = _spec_of(self._conv)
        self._frame = same_frame(self._conv)
        self._up = 2 if up_scale else 1
        self._noise_idx = self._conv_idx + 1 if add_noise else None
        self._norm_idx = (self._conv_idx + 1 + int(add_noise)) if norm_layer is not None else None
        if norm_layer is not None and not isinstance(self.de_conv_block[self._norm_idx], nn.BatchNorm2d):
            raise NotImplementedError("DeConvBlock on the CUDA path supports norm_layer in (None, nn.BatchNorm2d)")

    @property
    def _conv(self):
        return self.de_conv_block[self._conv_idx]

    def forward(self, x, seg=None):
        plain = not isinstance(x, Fn.Nhwc)
        xa = Fn.to_nhwc(x)
        conv = self._conv
        if self._simple and self._frame is None:  # DefectGAN heads
            y = Fn.conv(xa, conv, self._spec, conv.training, bias=conv.bias, act=self._act, out_fp32=True)
            return _edge(y) if plain else y
        # WGanGenerator blocks (generator.py:18-30): nearest 2x up-sampling and the 'same' zero frame in one pass, 'valid'
        # convolution with the noise fused into its epilogue, BatchNorm + activation in one pass
        if self._up > 1 or self._frame is not None:
            xa = Fn.upsample_pad(xa, self._up, self._frame or (0, 0, 0, 0))
        n, _, h, w = xa.shape
        ho, wo = self._spec.out_hw(h, w)
        noise, nw = _noise_args(self.de_conv_block[self._noise_idx] if self._noise_idx is not None else None, n, ho, wo,
                                xa.t.device)
        if self._norm_idx is None:
            fp32 = self._act in ("tanh", "sigmoid")
            y = Fn.conv(xa, conv, self._spec, conv.training, bias=conv.bias, noise=noise, noise_w=nw, act=self._act, out_fp32=fp32)
            if fp32 and not plain:
                y = Fn.Nhwc(y.t.to(Fn.ops.ACT_DTYPE), y.c)
        else:
            y = Fn.conv(xa, conv, self._spec, conv.training, bias=conv.bias, noise=noise, noise_w=nw)
            y = Fn.batch_norm(y, self.de_conv_block[self._norm_idx], act=self._act)
        return _edge(y) if plain else y


class ResBlock(nn.Module):
    """architecture.py:121-176: shortcut(x) + ConvBlock(act=None)([AvgPool2d] ConvBlock(act)(x)).  DefectGAN's encoder uses it with
    BatchNorm and the identity shortcut (generator.py:118-126: the add is fused into the second BatchNorm kernel); the
    StyleExtractor with InstanceNorm2d and `down_scale=True` (extractor.py:60-69: AvgPool2d(2, 2) between the two blocks and a
    1x1 ConvBlock + AvgPool2d shortcut)."""

    def __init__(self, f_in, f_out, kernel_size=(3, 3), stride=(1, 1), padding=0, padding_mode="zeros", bias=False,
                 norm_layer=nn.InstanceNorm2d, act_layer="relu", use_spectral=False, down_scale=False):
        super().__init__()
        if norm_layer is None:
            raise NotImplementedError("ResBlock without a norm layer")
        self.down_scale = down_scale
        blocks = [ConvBlock(f_in, f_in, kernel_size=kernel_size, stride=stride, padding=padding, padding_mode=padding_mode,
                            bias=bias, norm_layer=norm_layer, act_layer=act_layer, use_spectral=use_spectral),
                  ConvBlock(f_in, f_out, kernel_size=kernel_size, stride=stride, padding=padding, padding_mode=padding_mode,
                            bias=bias, norm_layer=norm_layer, act_layer=None, use_spectral=use_spectral)]
        if down_scale:
            blocks.insert(1, nn.AvgPool2d(2, 2))
            self.conv_s = nn.Sequential(ConvBlock(f_in, f_out, kernel_size=(1, 1), stride=(1, 1), padding=0, padding_mode=padding_mode,
                                                  bias=False, norm_layer=norm_layer, act_layer=None, use_spectral=use_spectral),
                                        nn.AvgPool2d(2, 2))
        self.res_block = nn.Sequential(*blocks)

    def forward(self, x, seg=None):
        plain = not isinstance(x, Fn.Nhwc)
        xa = Fn.to_nhwc(x)
        first, last = self.res_block[0], self.res_block[-1]
        if self.down_scale:
            y = Fn.avg_pool(first(xa), 2)
            y = last(y, residual=Fn.avg_pool(self.conv_s[0](xa), 2))
        else:
            y = first(xa, out_pad=last.reflect_pad_in())
            y = last(y, residual=xa)
        return Fn.from_nhwc(y) if plain else y


def next_pad(blocks, i) -> int:
    """Reflect padding the block after blocks[i] wants pre-written on its input (0 at the end of the chain or when the next
    block is not a plain ConvBlock, e.g. a ResBlock that also uses its input as the residual)."""
    nxt = blocks[i + 1] if i + 1 < len(blocks) else None
    return nxt.reflect_pad_in() if isinstance(nxt, ConvBlock) else 0


def _fusable_pad(spec: ConvSpec) -> int:
    """Reflect padding a producer may write for this convolution (0: zero padding / none)."""
    return spec.pad if (FUSE_REFLECT_PAD and spec.pad_mode == "reflect" and spec.pad > 0) else 0


def _consumer_spec(spec: ConvSpec, x: "Fn.Nhwc") -> ConvSpec:
    """The convolution as seen by an input that already carries its reflect padding (x.pad == spec.pad) or not (x.pad == 0)."""
    if x.pad:
        assert x.pad == spec.pad and spec.pad_mode == "reflect"
        return spec.unpadded()
    return spec


def _style_norm(kind, label_nc, nc, hidden_nc, kernel_size, padding, norm_layer, embed_nc=None, style_distill=False):
    """architecture.py:211-226 / 297-320: the style normalisation of a NormConvBlock / NormResBlock."""
    if kind == "spade":
        return SPADE(label_nc, nc, hidden_nc=hidden_nc, kernel_size=kernel_size, padding=padding, norm_layer=norm_layer)
    if kind == "sean":
        assert embed_nc is not None, "embed_nc must be specified for SEAN"
        return SEAN(embed_nc, nc, label_nc, hidden_nc=hidden_nc, norm_layer=norm_layer, style_distill=style_distill)
    if kind == "adain":
        return AdaIN(nc, hidden_nc=hidden_nc, norm_layer=norm_layer, denorm_type="linear")
    raise ValueError(f"Unknown style norm block type: {kind}")


def _style_forward(kind, norm, x, labels, style_feat, up, relu, out_pad):
    """norm_forward of architecture.py:246-253 / 362-369 with the fused followers (2x up-sampling, ReLU, reflect border)."""
    if kind == "spade":
        return norm(x, labels, up=up, relu=relu, out_pad=out_pad)
    if isinstance(labels, Fn.Nhwc):  # the generator hands SPADE a carrier; the vector norms want the [N, label_nc] rows
        labels = labels.t[:, 0, 0, : labels.c].float()
    if kind == "sean":
        return norm(x, labels, style_feat, up=up, relu=relu)
    return norm(x, style_feat, up=up, relu=relu)


class NormConvBlock(nn.Module):
    """architecture.py:179-257: noise(conv(act(SPADE(up2x(x))))) - the 2x nearest up-sampling and the ReLU are fused
    into the SPADE modulation kernel, the noise into the conv epilogue."""

    def __init__(self, style_norm_block_type, hidden_nc, label_nc, f_in, f_out, style_distill=False, embed_nc=None,
                 kernel_size=(3, 3), stride=(1, 1), padding=0, padding_mode="zeros", bias=False, up_scale=False,
                 norm_layer=nn.InstanceNorm2d, act_layer="relu", use_spectral=False, add_noise=False):
        super().__init__()
        self.up = nn.Upsample(scale_factor=2) if up_scale else nn.Identity()
        self.noise = NoiseInjection() if add_noise else nn.Identity()
        self.style_norm_block_type = style_norm_block_type
        self.norm = _style_norm(style_norm_block_type, label_nc, f_in, hidden_nc, (3, 3), padding, norm_layer, embed_nc, style_distill)
        self.conv = nn.Conv2d(f_in, f_out, kernel_size=kernel_size, stride=stride, padding=padding,
                              padding_mode=padding_mode, bias=bias)
        self.act = get_act_layer(act_layer)
        if use_spectral:
            self.conv = spectral_norm(self.conv)
        self._spec = _spec_of(self.conv)
        self._up = 2 if up_scale else 1
        if act_layer not in ("relu", None):
            raise NotImplementedError("NormConvBlock on the CUDA path fuses ReLU (or no activation) into the SPADE kernel")
        self._relu = act_layer == "relu"

    def forward(self, x, labels, style_feat=None, stats_out=False, out_pad=0):
        """stats_out: the output feeds another SPADE block - its per-image statistics come out of the convolution epilogue;
        out_pad: reflect padding the next convolution (the generator heads) wants pre-written on the output."""
        plain = not isinstance(x, Fn.Nhwc)
        xa = Fn.to_nhwc(x)
        s = _style_forward(self.style_norm_block_type, self.norm, xa, labels, style_feat, self._up, self._relu, _fusable_pad(self._spec))
        n, _, h, w = s.shape
        ho, wo = self._spec.out_hw(h, w)
        noise, nw = _noise_args(self.noise, n, ho, wo, s.t.device)
        y = Fn.conv(s, self.conv, _consumer_spec(self._spec, s), self.conv.training, bias=self.conv.bias, noise=noise, noise_w=nw,
                    want_stats="sample" if (stats_out and not plain) else None, out_pad=0 if plain else out_pad)
        return Fn.from_nhwc(y) if plain else y

    def update_alpha(self, epoch, num_epochs):
        if hasattr(self.norm, "update_alpha"):
            self.norm.update_alpha(epoch, num_epochs)


class NormResBlock(nn.Module):
    """architecture.py:260-371 with up_scale=False (generator.py:143-155): identity shortcut; ``norm_s`` / ``conv_s``
    are constructed (state-dict parity) but, exactly as in the reference, never executed."""

    def __init__(self, style_norm_block_type, hidden_nc, label_nc, f_in, f_out, style_distill=False, embed_nc=None,
                 kernel_size=(3, 3), stride=(1, 1), padding=0, padding_mode="zeros", bias=False, up_scale=False,
                 norm_layer=nn.InstanceNorm2d, act_layer="relu", use_spectral=False, add_noise=False):
        super().__init__()
        if up_scale:
            raise NotImplementedError("NormResBlock(up_scale=True) is not used by DefectGAN and not on the CUDA path")
        self.up = nn.Upsample(scale_factor=2)
        self.up_scale = up_scale
        self.noise_0 = NoiseInjection() if add_noise else nn.Identity()
        self.noise_1 = NoiseInjection() if add_noise else nn.Identity()
        f_mid = min(f_in, f_out)
        self.style_norm_block_type = style_norm_block_type
        self.norm_0 = _style_norm(style_norm_block_type, label_nc, f_in, hidden_nc, kernel_size, padding, norm_layer, embed_nc, style_distill)
        self.norm_1 = _style_norm(style_norm_block_type, label_nc, f_mid, hidden_nc, kernel_size, padding, norm_layer, embed_nc, style_distill)
        self.norm_s = _style_norm(style_norm_block_type, label_nc, f_in, hidden_nc, kernel_size, padding, norm_layer, embed_nc, style_distill)
        self.act = get_act_layer(act_layer)
        mk = lambda a, b: nn.Conv2d(a, b, kernel_size=kernel_size, stride=stride, padding=padding,  # noqa: E731
                                    padding_mode=padding_mode, bias=bias)
        self.conv_0, self.conv_1, self.conv_s = mk(f_in, f_mid), mk(f_mid, f_out), mk(f_in, f_out)
        if use_spectral:
            self.conv_0 = spectral_norm(self.conv_0)
            self.conv_1 = spectral_norm(self.conv_1)
            self.conv_s = spectral_norm(self.conv_s)
        self._spec0, self._spec1 = _spec_of(self.conv_0), _spec_of(self.conv_1)
        if act_layer not in ("relu", None):
            raise NotImplementedError("NormResBlock on the CUDA path fuses ReLU (or no activation) into the SPADE kernel")
        self._relu = act_layer == "relu"

    def forward(self, x, labels, style_feat=None, stats_out=False):
        """stats_out: the output feeds another SPADE block (its per-image statistics then come out of the convolution epilogue)."""
        plain = not isinstance(x, Fn.Nhwc)
        xa = Fn.to_nhwc(x)
        n, _, h, w = xa.shape
        dev = xa.t.device
        kind = self.style_norm_block_type
        s = _style_forward(kind, self.norm_0, xa, labels, style_feat, 1, self._relu, _fusable_pad(self._spec0))
        noise, nw = _noise_args(self.noise_0, n, h, w, dev)
        y = Fn.conv(s, self.conv_0, _consumer_spec(self._spec0, s), self.conv_0.training, bias=self.conv_0.bias, noise=noise, noise_w=nw,
                    want_stats="sample")
        s = _style_forward(kind, self.norm_1, y, labels, style_feat, 1, self._relu, _fusable_pad(self._spec1))
        noise, nw = _noise_args(self.noise_1, n, h, w, dev)
        y = Fn.conv(s, self.conv_1, _consumer_spec(self._spec1, s), self.conv_1.training, bias=self.conv_1.bias, noise=noise, noise_w=nw,
                    residual=xa, want_stats="sample" if (stats_out and not plain) else None)
        return Fn.from_nhwc(y) if plain else y

    def update_alpha(self, epoch, num_epochs):
        pass


class _MaskBlend(torch.autograd.Function):
    """masked = imgs * m + token * (1 - m); differentiable in the (learnable) token only - the images are data."""

    @staticmethod
    def forward(ctx, imgs, token, coarse, hs, ws, patch):
        out, mask = Fn.ops.mask_token_fwd(imgs, coarse, hs, ws, patch, token=token)
        ctx.geo = (hs, ws, patch, tuple(token.shape) if token is not None else None)
        ctx.save_for_backward(coarse)
        ctx.mark_non_differentiable(mask)
        return out, mask

    @staticmethod
    def backward(ctx, dy, _dmask):
        (coarse,) = ctx.saved_tensors
        hs, ws, patch, tshape = ctx.geo
        dtok = None
        if tshape is not None and ctx.needs_input_grad[1]:
            dtok = Fn.ops.mask_token_bwd(dy.contiguous().float(), coarse, hs, ws, patch, tshape)
        return None, dtok, None, None, None, None


class MaskToken(nn.Module):
    """architecture.py:392-418: replaces the masked-out patches by a token - zero, the per-image mean of the visible pixels
    ('mean'), or a learnable scalar / per-channel vector / per-position map / full image.  Same parameter name and shapes (state-dict
    key ``mask_token``).  ``forward(imgs, masks)`` is the reference's formula on a materialised mask (drop-in use);
    ``blend(imgs, coarse, h_shift, w_shift, patch_size)`` is the fused path used by DefectGanModel._repair_mask: the shifted patch mask
    is evaluated on the fly from the coarse Bernoulli grid (utils.util.draw_shifted_mask) by dei2i_mask_token_fwd."""

    def __init__(self, opt):
        super().__init__()
        self.mask_token_type = opt.mask_token_type
        self.mask_ratio = opt.mask_ratio
        if self.mask_token_type in ("zero", "mean"):
            self.mask_token = 0
        elif self.mask_token_type == "scalar":
            self.mask_token = nn.Parameter(torch.zeros(1, 1, 1, 1))
        elif self.mask_token_type == "vector":
            self.mask_token = nn.Parameter(torch.zeros(1, opt.input_nc, 1, 1))
        elif self.mask_token_type == "position":
            self.mask_token = nn.Parameter(torch.zeros(1, 1, opt.image_size, opt.image_size))
        elif self.mask_token_type == "full":
            self.mask_token = nn.Parameter(torch.zeros(1, opt.input_nc, opt.image_size, opt.image_size))
        else:
            raise ValueError("Unknown mask token type: {}".format(self.mask_token_type))

    def forward(self, imgs, masks):
        masked_imgs = imgs * masks
        if self.mask_token_type == "mean":
            self._calc_mean_mask_token(masked_imgs)
        return masked_imgs + self.mask_token * (1 - masks)

    def _calc_mean_mask_token(self, masked_imgs):
        img_mean = masked_imgs.mean(dim=(2, 3)) / self.mask_ratio
        self.mask_token = img_mean.reshape(*img_mean.size()[:2], 1, 1)

    def blend(self, imgs, coarse, h_shift, w_shift, patch_size):
        """-> (masked_imgs, masks [B,1,H,W]) on the fused kernels."""
        imgs = imgs.contiguous().float()
        coarse = coarse.to(imgs.device, non_blocking=True).contiguous().float()
        token = None
        if self.mask_token_type == "mean":
            token = Fn.ops.mask_mean(imgs, coarse, h_shift, w_shift, patch_size, self.mask_ratio)
            self.mask_token = token  # the reference keeps the last per-image mean on the module too (architecture.py:416-418)
        elif isinstance(self.mask_token, nn.Parameter):
            token = self.mask_token
        return _MaskBlend.apply(imgs, token, coarse, int(h_shift), int(w_shift), int(patch_size))


class EmbedEncoder(nn.Module):
    """architecture.py:420-431 (imported by extractor.py; only the commented-out StyleExtractor used it)."""

    def __init__(self, embed_nc, hidden_nc):
        super().__init__()
        self.mlp_shared = nn.Sequential(nn.Linear(embed_nc, hidden_nc), nn.ReLU(inplace=True),
                                        nn.Linear(hidden_nc, hidden_nc), nn.ReLU(inplace=True))

    def forward(self, feat):
        if feat.dim() == 3:
            feat = feat.mean(dim=1)
        h = Fn.linear(feat.float(), self.mlp_shared[0], act="relu")
        return Fn.linear(h, self.mlp_shared[2], act="relu")


class LatentDecoder(nn.Module):
    """architecture.py:434-448: labels | N(0, 1) noise -> two Linear + ReLU.  The noise is drawn on the host exactly as the
    reference draws it (`torch.randn(...).to(device)`), so a seeded run sees the same values."""

    def __init__(self, label_nc, hidden_nc, latent_dim):
        super().__init__()
        self.mlp_latent = nn.Sequential(nn.Linear(latent_dim, hidden_nc // 2), nn.ReLU(inplace=True),
                                        nn.Linear(hidden_nc // 2, hidden_nc), nn.ReLU(inplace=True))
        self.noise_dim = latent_dim - label_nc

    def forward(self, labels):
        if labels.dim() == 4:
            labels = labels.view(labels.size(0), -1)
        noise = torch.randn(labels.size(0), self.noise_dim).to(labels.device)
        latent = torch.cat([labels.float(), noise], dim=1)
        h = Fn.linear(latent, self.mlp_latent[0], act="relu")
        return Fn.linear(h, self.mlp_latent[2], act="relu")
