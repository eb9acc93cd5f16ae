"""DefectGanDiscriminator (models/networks/discriminator.py:49-98) on the sm_100a kernels."""
import numpy as np
import torch
from torch import nn

from ... import functional as Fn
from .architecture import ConvBlock, next_pad
from .base_network import BaseNetwork


class WGanDiscriminator(BaseNetwork):
    """models/networks/discriminator.py:10-46: 7x7 s2 reflect conv + BN + ReLU, MaxPool 3/2/1, [3x3 s2 conv + BN + ReLU] x
    num_layers, global average pool, Linear -> logits [B, 1]."""

    def __init__(self, num_layers, ndf=64):
        super().__init__()
        dim = ndf
        stem = [ConvBlock(3, ndf, kernel_size=(7, 7), stride=(2, 2), padding=3, padding_mode="reflect", norm_layer=nn.BatchNorm2d,
                          act_layer="relu"),
                nn.MaxPool2d(kernel_size=(3, 3), stride=(2, 2), padding=1)]
        backbone = []
        for _ in range(num_layers):
            backbone += [ConvBlock(dim, dim * 2, kernel_size=(3, 3), stride=(2, 2), padding=1, bias=False, norm_layer=nn.BatchNorm2d,
                                   act_layer="relu")]
            dim *= 2
        backbone += [nn.AdaptiveAvgPool2d(1)]
        self.feature_extractor = nn.Sequential(*(stem + backbone))
        self.classifier = nn.Sequential(nn.Linear(dim, 1))

    def forward(self, x):
        x = x.to(self.device)
        feat = Fn.to_nhwc(x)
        pooled = None
        for m in self.feature_extractor:
            if isinstance(m, ConvBlock):
                feat = m(feat)
            elif isinstance(m, nn.MaxPool2d):
                k, s_, p_ = (v[0] if isinstance(v, (tuple, list)) else v for v in (m.kernel_size, m.stride, m.padding))
                feat = Fn.max_pool(feat, k, s_, p_)
            else:  # AdaptiveAvgPool2d(1) + view(batch, -1)
                pooled = Fn.global_avg_pool(feat)
        return Fn.linear(pooled, self.classifier[0])


class DefectGanDiscriminator(BaseNetwork):
    def __init__(self, opt):
        super().__init__()
        dim = opt.ndf
        common = dict(kernel_size=(4, 4), stride=(2, 2), padding=1, padding_mode="reflect", norm_layer=None,
                      act_layer="leaky_relu", use_spectral=opt.use_spectral)
        blocks = [ConvBlock(opt.input_nc, dim, **common)]
        for _ in range(opt.num_layers):
            blocks.append(ConvBlock(dim, dim * 2, **common))
            dim *= 2
        k = int(opt.image_size // np.power(2, opt.num_layers + 1))
        self.enc_blk = nn.Sequential(*blocks)
        self.cls_clf = ConvBlock(dim, opt.label_nc, kernel_size=(k, k), norm_layer=None, act_layer=None)
        self.src_clf = ConvBlock(dim, 1, kernel_size=(3, 3), stride=(1, 1), padding="same", padding_mode="reflect",
                                 norm_layer=None, act_layer=None)

    def forward(self, x):
        """(src_logits [B,1,h,h], cls_logits [B,label_nc]) like discriminator.py:92-98."""
        assert isinstance(x, torch.Tensor), "x must be Original Images: Torch.Tensor"
        feat = Fn.to_nhwc(x)
        Fn.prefetch_spectral_norm(self, self.training)
        Fn.prefetch_packs(self)
        enc = list(self.enc_blk)
        for i, blk in enumerate(enc):
            feat = blk(feat, out_pad=next_pad(enc, i)) if isinstance(blk, ConvBlock) else blk(feat)
        src = self._logits(self.src_clf, feat)
        cls = self._logits(self.cls_clf, feat)
        return src, cls.reshape(cls.size(0), cls.size(1))

    @staticmethod
    def _logits(block: ConvBlock, feat: Fn.Nhwc) -> torch.Tensor:
        conv = block.conv_block[0]
        y = Fn.conv(feat, conv, block._spec, conv.training, bias=conv.bias, out_fp32=True)
        return y.t[..., : y.c].permute(0, 3, 1, 2).contiguous()
