"""WGanTrainer (trainers/wgan_trainer.py:17-138): the two step functions of the weight-clipping WGAN.

The reference class cannot be constructed as checked in (``BaseTrainer._init_lr`` indexes the scalar ``--lr`` of
options/wgan_options.py:52); the mirror accepts a scalar or a list and otherwise keeps the reference's step
bodies line for line: weight clipping -> G(B) with uniform noise -> D(fake.detach()), D(real) -> -(E[real] - E[fake]) ->
RMSprop; generator: -E[D(G(z))]."""
import math

import torch

from .. import functional as Fn
from .base_trainer import BaseTrainer


class WGanTrainer(BaseTrainer):
    def __init__(self, opt, iters_per_epoch=math.inf):
        if not hasattr(opt, "iters_per_epoch"):
            opt.iters_per_epoch = iters_per_epoch
        super().__init__(opt)
        self.fix_noise = torch.rand(opt.num_display_images, opt.noise_dim, 1, 1)

    def _init_lr(self, opt):
        self.lr = opt.lr[0] if isinstance(opt.lr, (list, tuple)) else opt.lr

    @torch.no_grad()
    def _generate_image(self):
        fake = self.model.netG(self.fix_noise)
        return (fake / 2 + 0.5).detach().cpu()

    def _train_generator_once(self, batch_size):
        """wgan_trainer.py:115-122.  ``batch_size`` may also be a noise tensor (netG accepts both, generator.py:41-47)."""
        d_params = list(self.model.netD.parameters())
        for p in d_params:  # the reference computes the critic's gradients here and discards them (D zero_grad precedes their use)
            p.requires_grad_(False)
        try:
            with Fn.weight_cache():
                self.optimizers["G"].zero_grad()
                fake_data = self.model.netG(batch_size)
                fake_logits = self.model.netD(fake_data)
                g_loss = -fake_logits.mean()
                g_loss.backward()
        finally:
            for p in d_params:
                p.requires_grad_(True)
        self._reduce_and_step("G")
        self.losses["gan_G"].append(g_loss.item())

    def _train_discriminator_once(self, real_data, noise=None):
        """wgan_trainer.py:124-138."""
        self.model.weight_clipping()
        with Fn.weight_cache():
            self.optimizers["D"].zero_grad()
            with torch.no_grad():
                fake_data = self.model.netG(real_data.shape[0] if noise is None else noise)
            fake_logits = self.model.netD(fake_data.detach_())
            real_logits = self.model.netD(real_data)
            d_loss = -(real_logits.mean() - fake_logits.mean())
            d_loss.backward()
        self._reduce_and_step("D")
        self.losses["gan_D"].append(d_loss.item())
        self.dis_outputs["real"] += real_logits.flatten().tolist()
        self.dis_outputs["fake"] += fake_logits.flatten().tolist()
