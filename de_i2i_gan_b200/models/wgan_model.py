"""WGanModel (models/wgan_model.py:8-21): generator + critic of the weight-clipping WGAN on the sm_100a kernels."""
import math

import torch

from .. import ops
from .base_model import BaseModel
from .networks.discriminator import WGanDiscriminator
from .networks.generator import WGanGenerator


class WGanModel(BaseModel):
    def __init__(self, opt):
        super().__init__(opt)
        image_size = opt.image_size
        assert image_size & (image_size - 1) == 0, "Image size must be a power of 2"
        num_layers = int(math.log(image_size, 2)) - 3
        self.netG = WGanGenerator(num_layers=num_layers, ngf=opt.ngf, noise_dim=opt.noise_dim).to(opt.device)
        self.netD = WGanDiscriminator(num_layers=num_layers, ndf=opt.ndf).to(opt.device)
        self.clipping_limit = opt.clipping_limit

    @torch.no_grad()
    def weight_clipping(self):
        """wgan_model.py:18-21: clamp every critic parameter to [-c, c] (dei2i_clamp, in place)."""
        for p in self.netD.parameters():
            ops.clamp_(p.data, -self.clipping_limit, self.clipping_limit)
