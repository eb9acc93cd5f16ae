"""StyleExtractor (models/networks/extractor.py:36-96) on the sm_100a kernels: the style vector of the AdaIN variant of DefectGAN.

sean_alpha == 1: 7x7 stride-2 reflect ConvBlock + LeakyReLU, log2(image_size) - 3 down-scaling ResBlocks (InstanceNorm2d, LeakyReLU,
AvgPool2d(2, 2)), 4x4 valid ConvBlock -> [N, hidden_nc, 1, 1]; sean_alpha == 0: labels | noise -> five Linear layers.
Module tree and state-dict keys are the reference's (`shared.*`).
"""
import math

import torch
from torch import nn

from ... import functional as Fn
from .architecture import ConvBlock, ResBlock
from .base_network import BaseNetwork


class StyleExtractor(BaseNetwork):
    def __init__(self, opt):
        super().__init__()
        assert opt.image_size in (64, 128, 256, 512, 1024), "image size should be one of [64, 128, 256, 512, 1024]"
        num_blocks = int(math.log2(opt.image_size)) - 3
        crt_dim = opt.ndf
        max_dim = 256
        self.sean_alpha = opt.sean_alpha
        self.noise_dim = opt.latent_dim - opt.label_nc
        if opt.sean_alpha == 0:
            layers = [nn.Linear(opt.latent_dim, max_dim), nn.ReLU(inplace=True)]
            for _ in range(3):
                layers += [nn.Linear(max_dim, max_dim), nn.ReLU(inplace=True)]
            layers.append(nn.Linear(max_dim, opt.hidden_nc))
            self.shared = nn.Sequential(*layers)
        elif opt.sean_alpha == 1:
            blocks = [ConvBlock(opt.input_nc, crt_dim, kernel_size=(7, 7), stride=(2, 2), padding=3, padding_mode="reflect",
                                norm_layer=None, act_layer="leaky_relu", use_spectral=False)]
            for _ in range(num_blocks):
                new_dim = min(crt_dim * 2, max_dim)
                blocks.append(ResBlock(crt_dim, new_dim, kernel_size=(3, 3), stride=(1, 1), padding="same", padding_mode="reflect",
                                       norm_layer=nn.InstanceNorm2d, act_layer="leaky_relu", use_spectral=False, down_scale=True))
                crt_dim = new_dim
            blocks.append(ConvBlock(crt_dim, opt.hidden_nc, kernel_size=(4, 4), stride=(1, 1), padding=0, norm_layer=None,
                                    act_layer=None, use_spectral=False))
            self.shared = nn.Sequential(*blocks)
        else:
            raise NotImplementedError("sean_alpha should be 0 or 1")

    def forward(self, x, labels):
        if self.sean_alpha == 0:
            # host draw, as extractor.py:85 (`torch.randn(...).to(x.device)`): a seeded run sees the reference's values
            noise = torch.randn(labels.size(0), self.noise_dim).to(x.device)
            h = torch.cat([labels.reshape(labels.size(0), -1).float(), noise], dim=1)
            mods = list(self.shared)
            for i, m in enumerate(mods):
                if isinstance(m, nn.Linear):
                    relu = i + 1 < len(mods) and isinstance(mods[i + 1], nn.ReLU)
                    h = Fn.linear(h, m, act="relu" if relu else None)
            return h
        blocks = list(self.shared)
        feat = blocks[0](x, carrier_out=True)
        for blk in blocks[1:]:
            feat = blk(feat)
        return Fn.from_nhwc(feat)
