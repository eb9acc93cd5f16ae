"""StarGAN-v2 networks (stargan-v2/core/model.py:26-123, 321-393, 442-532; ``norm_type='adain'``, ``w_hpf=0``) on the sm_100a
kernels.  Same class names, constructor signatures, sub-module names and registration order as the reference, so that
``state_dict()`` keys and initialiser RNG order are identical and reference checkpoints load unchanged.  ``nn.Conv2d`` /
``nn.Linear`` / ``nn.InstanceNorm2d`` objects are parameter containers; every ``forward`` runs the CUDA path:

* convolutions: ``Fn.conv_linear`` (TMA / tcgen05 implicit GEMM, bias in the epilogue), twice differentiable along the
  input-gradient path - the discriminator's R1 penalty (solver.py:573-583) is a gradient of a gradient;
* InstanceNorm(affine) / AdaIN + LeakyReLU(0.2): one fused kernel (``Fn.adain_act``);
* pre-activation LeakyReLU, 2x average pooling, nearest 2x up-sampling, (x + r) / sqrt(2): streaming kernels.

Not mirrored (outside the hot path): the SEAN blocks, ``HighPass`` / FAN masks (``w_hpf > 0``), the ViT
``FeatureExtractor`` and ``build_model`` (needs downloaded weights).
"""
import copy
import math

import numpy as np
import torch
import torch.nn as nn

from ... import functional as Fn

_SQRT1_2 = 1.0 / math.sqrt(2.0)


def _nc(norm: nn.InstanceNorm2d, n: int):
    """Per-sample (scale, shift) of an affine InstanceNorm2d: weight / bias broadcast over the batch."""
    return norm.weight.unsqueeze(0).expand(n, -1), norm.bias.unsqueeze(0).expand(n, -1)


class ResBlk(nn.Module):
    """model.py:26-68 (pre-activation residual block; optional affine IN, optional 2x average-pool down-sampling)."""

    def __init__(self, dim_in, dim_out, actv=nn.LeakyReLU(0.2), normalize=False, downsample=False):
        super().__init__()
        self.actv = actv
        self.normalize = normalize
        self.downsample = downsample
        self.learned_sc = dim_in != dim_out
        self._build_weights(dim_in, dim_out)
        self._slope = float(getattr(actv, "negative_slope", 0.2))

    def _build_weights(self, dim_in, dim_out):
        self.conv1 = nn.Conv2d(dim_in, dim_in, 3, 1, 1)
        self.conv2 = nn.Conv2d(dim_in, dim_out, 3, 1, 1)
        if self.normalize:
            self.norm1 = nn.InstanceNorm2d(dim_in, affine=True)
            self.norm2 = nn.InstanceNorm2d(dim_in, affine=True)
        if self.learned_sc:
            self.conv1x1 = nn.Conv2d(dim_in, dim_out, 1, 1, 0, bias=False)

    def _shortcut(self, x):
        if self.learned_sc:
            x = Fn.conv_linear(x, self.conv1x1)
        if self.downsample:
            x = Fn.avg_pool(x, 2)
        return x

    def _norm_act(self, x, norm):
        if self.normalize:
            return Fn.adain_act(x, *_nc(norm, x.t.shape[0]), slope=self._slope, eps=norm.eps)
        return Fn.activation(x, "leaky_relu")

    def _residual(self, x):
        x = self._norm_act(x, getattr(self, "norm1", None))
        x = Fn.conv_linear(x, self.conv1)
        if self.downsample:
            x = Fn.avg_pool(x, 2)
        x = self._norm_act(x, getattr(self, "norm2", None))
        return Fn.conv_linear(x, self.conv2)

    def forward(self, x):
        plain = not isinstance(x, Fn.Nhwc)
        xa = Fn.to_nhwc(x)
        y = Fn.add_scaled(self._shortcut(xa), self._residual(xa), _SQRT1_2)
        return Fn.from_nhwc(y) if plain else y


class AdaIN(nn.Module):
    """model.py:71-84: (1 + gamma(s)) * IN(x) + beta(s); on the CUDA path fused with the following LeakyReLU (``slope``)."""

    def __init__(self, style_dim, num_features):
        super().__init__()
        self.norm = nn.InstanceNorm2d(num_features, affine=False)
        self.fc = nn.Linear(style_dim, num_features * 2)

    def forward(self, x, s, slope=1.0):
        plain = not isinstance(x, Fn.Nhwc)
        xa = Fn.to_nhwc(x)
        h = Fn.linear(s.float(), self.fc)
        gamma, beta = torch.chunk(h, chunks=2, dim=1)
        y = Fn.adain_act(xa, 1 + gamma, beta, slope=slope, eps=self.norm.eps)
        return Fn.from_nhwc(y) if plain else y


class AdainResBlk(nn.Module):
    """model.py:87-123 with w_hpf = 0."""

    def __init__(self, dim_in, dim_out, style_dim=64, w_hpf=0, actv=nn.LeakyReLU(0.2), upsample=False):
        super().__init__()
        if w_hpf != 0:
            raise NotImplementedError("w_hpf > 0 (FAN heat-map high-pass skip connections) is outside the hot path")
        self.w_hpf = w_hpf
        self.actv = actv
        self.upsample = upsample
        self.learned_sc = dim_in != dim_out
        self._build_weights(dim_in, dim_out, style_dim)
        self._slope = float(getattr(actv, "negative_slope", 0.2))

    def _build_weights(self, dim_in, dim_out, style_dim=64):
        self.conv1 = nn.Conv2d(dim_in, dim_out, 3, 1, 1)
        self.conv2 = nn.Conv2d(dim_out, dim_out, 3, 1, 1)
        self.norm1 = AdaIN(style_dim, dim_in)
        self.norm2 = AdaIN(style_dim, dim_out)
        if self.learned_sc:
            self.conv1x1 = nn.Conv2d(dim_in, dim_out, 1, 1, 0, bias=False)

    def _shortcut(self, x):
        if self.upsample:
            x = Fn.upsample_pad(x, 2)
        if self.learned_sc:
            x = Fn.conv_linear(x, self.conv1x1)
        return x

    def _residual(self, x, s):
        x = self.norm1(x, s, slope=self._slope)
        if self.upsample:
            x = Fn.upsample_pad(x, 2)
        x = Fn.conv_linear(x, self.conv1)
        x = self.norm2(x, s, slope=self._slope)
        return Fn.conv_linear(x, self.conv2)

    def forward(self, x, s, labels=None):
        plain = not isinstance(x, Fn.Nhwc)
        xa = Fn.to_nhwc(x)
        y = Fn.add_scaled(self._residual(xa, s), self._shortcut(xa), _SQRT1_2)
        return Fn.from_nhwc(y) if plain else y


class Generator(nn.Module):
    """model.py:321-393 (norm_type 'adain')."""

    def __init__(self, img_size=256, style_dim=64, max_conv_dim=512, w_hpf=1, norm_type="adain", embed_nc=768, label_nc=3,
                 hidden_nc=256):
        super().__init__()
        if norm_type != "adain":
            raise NotImplementedError("only norm_type='adain' is on the CUDA hot path (SEAN: SURVEY.md 8f, next)")
        dim_in = 2 ** 14 // img_size
        self.img_size = img_size
        self.from_rgb = nn.Conv2d(3, dim_in, 3, 1, 1)
        self.encode = nn.ModuleList()
        self.decode = nn.ModuleList()
        self.to_rgb = nn.Sequential(nn.InstanceNorm2d(dim_in, affine=True), nn.LeakyReLU(0.2), nn.Conv2d(dim_in, 3, 1, 1, 0))
        repeat_num = int(np.log2(img_size)) - 4
        if w_hpf > 0:
            repeat_num += 1
        dim_out = dim_in
        for _ in range(repeat_num):
            dim_out = min(dim_in * 2, max_conv_dim)
            self.encode.append(ResBlk(dim_in, dim_out, normalize=True, downsample=True))
            self.decode.insert(0, AdainResBlk(dim_out, dim_in, style_dim, w_hpf=w_hpf, upsample=True))  # stack-like
            dim_in = dim_out
        for _ in range(2):
            self.encode.append(ResBlk(dim_out, dim_out, normalize=True))
            self.decode.insert(0, AdainResBlk(dim_out, dim_out, style_dim, w_hpf=w_hpf))
        self.decoder_len = len(self.decode)

    def forward(self, x, s, masks=None, labels=None, layer_split_index=None):
        assert masks is None and layer_split_index is None, "masks / layer_split_index belong to the w_hpf > 0 and SEAN variants"
        h = Fn.conv_linear(Fn.to_nhwc(x), self.from_rgb)
        for block in self.encode:
            h = block(h)
        for block in self.decode:
            h = block(h, s, labels)
        norm, act, conv = self.to_rgb
        h = Fn.adain_act(h, *_nc(norm, h.t.shape[0]), slope=act.negative_slope, eps=norm.eps)
        y = Fn.conv_linear(h, conv, out_fp32=True)
        return y.t[..., : y.c].permute(0, 3, 1, 2).contiguous()


class MappingNetwork(nn.Module):
    """model.py:442-471."""

    def __init__(self, latent_dim=16, style_dim=64, num_domains=2):
        super().__init__()
        layers = [nn.Linear(latent_dim, 512), nn.ReLU()]
        for _ in range(3):
            layers += [nn.Linear(512, 512), nn.ReLU()]
        self.shared = nn.Sequential(*layers)
        self.unshared = nn.ModuleList()
        for _ in range(num_domains):
            self.unshared += [nn.Sequential(nn.Linear(512, 512), nn.ReLU(), nn.Linear(512, 512), nn.ReLU(), nn.Linear(512, 512),
                                            nn.ReLU(), nn.Linear(512, style_dim))]

    @staticmethod
    def _mlp(seq, h):
        mods = list(seq)
        i = 0
        while i < len(mods):
            relu = i + 1 < len(mods) and isinstance(mods[i + 1], nn.ReLU)
            h = Fn.linear(h, mods[i], act="relu" if relu else None)
            i += 2 if relu else 1
        return h

    def forward(self, z, y):
        h = self._mlp(self.shared, z.float())
        out = torch.stack([self._mlp(layer, h) for layer in self.unshared], dim=1)  # (batch, num_domains, style_dim)
        idx = torch.arange(y.size(0), device=y.device)
        return out[idx, y]


def _trunk(blocks, x):
    """Shared body of StyleEncoder.shared / Discriminator.main: conv, ResBlk x n, LeakyReLU, 4x4 valid conv, LeakyReLU[, 1x1 conv]."""
    h = Fn.to_nhwc(x)
    mods = list(blocks)
    for i, m in enumerate(mods):
        if isinstance(m, nn.Conv2d):
            h = Fn.conv_linear(h, m, out_fp32=(i == len(mods) - 1))
        elif isinstance(m, ResBlk):
            h = m(h)
        elif isinstance(m, nn.LeakyReLU):
            h = Fn.activation(h, "leaky_relu")
        else:
            raise NotImplementedError(type(m))
    return h


class StyleEncoder(nn.Module):
    """model.py:474-503."""

    def __init__(self, img_size=256, style_dim=64, num_domains=2, max_conv_dim=512):
        super().__init__()
        dim_in = 2 ** 14 // img_size
        blocks = [nn.Conv2d(3, dim_in, 3, 1, 1)]
        repeat_num = int(np.log2(img_size)) - 2
        dim_out = dim_in
        for _ in range(repeat_num):
            dim_out = min(dim_in * 2, max_conv_dim)
            blocks += [ResBlk(dim_in, dim_out, downsample=True)]
            dim_in = dim_out
        blocks += [nn.LeakyReLU(0.2), nn.Conv2d(dim_out, dim_out, 4, 1, 0), nn.LeakyReLU(0.2)]
        self.shared = nn.Sequential(*blocks)
        self.unshared = nn.ModuleList()
        for _ in range(num_domains):
            self.unshared += [nn.Linear(dim_out, style_dim)]

    def forward(self, x, y):
        h = _trunk(self.shared, x)
        h = h.t[..., : h.c].reshape(h.t.shape[0], -1).float()
        out = torch.stack([Fn.linear(h, layer) for layer in self.unshared], dim=1)  # (batch, num_domains, style_dim)
        idx = torch.arange(y.size(0), device=y.device)
        return out[idx, y]


class Discriminator(nn.Module):
    """model.py:506-532."""

    def __init__(self, img_size=256, num_domains=2, max_conv_dim=512):
        super().__init__()
        dim_in = 2 ** 14 // img_size
        blocks = [nn.Conv2d(3, dim_in, 3, 1, 1)]
        repeat_num = int(np.log2(img_size)) - 2
        dim_out = dim_in
        for _ in range(repeat_num):
            dim_out = min(dim_in * 2, max_conv_dim)
            blocks += [ResBlk(dim_in, dim_out, downsample=True)]
            dim_in = dim_out
        blocks += [nn.LeakyReLU(0.2), nn.Conv2d(dim_out, dim_out, 4, 1, 0), nn.LeakyReLU(0.2), nn.Conv2d(dim_out, num_domains, 1, 1, 0)]
        self.main = nn.Sequential(*blocks)

    def forward(self, x, y):
        h = _trunk(self.main, x)  # fp32 [N,1,1,num_domains (padded to 8)]
        out = h.t[..., : h.c].reshape(h.t.shape[0], -1)
        idx = torch.arange(y.size(0), device=y.device)
        return out[idx, y]


def build_nets(args):
    """The adain branch of build_model (model.py:695-717) without nn.DataParallel / FAN / ViT: (nets, nets_ema)."""
    from types import SimpleNamespace
    generator = Generator(args.img_size, args.style_dim, w_hpf=args.w_hpf, norm_type="adain")
    mapping_network = MappingNetwork(args.latent_dim, args.style_dim, args.num_domains)
    style_encoder = StyleEncoder(args.img_size, args.style_dim, args.num_domains)
    discriminator = Discriminator(args.img_size, args.num_domains)
    nets = SimpleNamespace(generator=generator, mapping_network=mapping_network, style_encoder=style_encoder,
                           discriminator=discriminator)
    nets_ema = SimpleNamespace(generator=copy.deepcopy(generator), mapping_network=copy.deepcopy(mapping_network),
                               style_encoder=copy.deepcopy(style_encoder))
    return nets, nets_ema
