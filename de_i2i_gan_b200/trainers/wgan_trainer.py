"""WGanTrainer (trainers/wgan_trainer.py:17-138): the two step functions of the weight-clipping WGAN.

The reference class cannot be constructed as checked in (``BaseTrainer._init_lr`` indexes the scalar ``--lr`` of
options/wgan_options.py:52); the mirror accepts a scalar or a list and otherwise keeps the reference's step
bodies line for line: weight clipping -> G(B) with uniform noise -> D(fake.detach()), D(real) -> -(E[real] - E[fake]) ->
RMSprop; generator: -E[D(G(z))]."""
import math
import os

import torch

from .. import functional as Fn
from .base_trainer import BaseTrainer
from .defectgan_trainer import GraphedStep


class WGanTrainer(BaseTrainer):
    def __init__(self, opt, iters_per_epoch=math.inf):
        if not hasattr(opt, "iters_per_epoch"):
            opt.iters_per_epoch = iters_per_epoch
        super().__init__(opt)
        self.fix_noise = torch.rand(opt.num_display_images, opt.noise_dim, 1, 1)
        # whole-step CUDA graphs (the step is ~290 small launches: 19 ms eager, host-bound).  The uniform generator noise is drawn on
        # the host with the reference's own call (generator.py:46) and enters the graph as an input.
        use_graph = getattr(opt, "cuda_graph", None)
        if use_graph is None:
            use_graph = os.environ.get("DEI2I_CUDA_GRAPH", "1") != "0"
        use_graph = bool(use_graph) and torch.device(opt.device).type == "cuda"
        pool = [None]
        self._graphed = {"D": GraphedStep(self._d_body, self.optimizers["D"], use_graph, pool),
                         "G": GraphedStep(self._g_body, self.optimizers["G"], use_graph, pool)}
        self._d_params = list(self.model.netD.parameters())

    def _init_lr(self, opt):
        self.lr = opt.lr[0] if isinstance(opt.lr, (list, tuple)) else opt.lr

    @torch.no_grad()
    def _generate_image(self):
        fake = self.model.netG(self.fix_noise)
        return (fake / 2 + 0.5).detach().cpu()

    # ------------------------------------------------------------------------------------------------ epoch loop
    def train(self, train_loader, val_loader=None):
        """wgan_trainer.py:69-84: epochs 1..num_epochs; per epoch the losses / D(x) records are cleared, the loader is walked
        (_train_epoch), the optional validation pass records D(x) on held-out images, TensorBoard scalars + an image grid of the
        fixed noise are written (only when ``opt.log_dir`` is set and tensorboard is importable) and a checkpoint is saved every
        ``save_ckpt_freq`` epochs."""
        writer = self._make_writer()
        for epoch in range(1, self.opt.num_epochs + 1):
            self.losses.clear()
            self.dis_outputs.clear()
            self._train_epoch(train_loader, epoch)
            if val_loader is not None:
                self._val_epoch(val_loader)
            if writer is not None:
                self._write_tf_log(writer, epoch)
            if epoch % self.opt.save_ckpt_freq == 0 and self.rank == 0:
                self.model.save(epoch)
        if writer is not None:
            writer.close()

    def _train_epoch(self, data_loader, epoch):
        """wgan_trainer.py:86-107 (progress bar omitted): one critic step per batch, one generator step every ``num_critics``
        iterations, the 'latest' checkpoint every ``save_latest_freq`` iterations, one scheduler step per epoch."""
        for batch_data in data_loader:
            self.iters += 1
            batch_data = batch_data.to(self.opt.device, non_blocking=True)
            self._train_discriminator_once(batch_data)
            if self.iters % self.opt.num_critics == 0:
                self._train_generator_once(batch_data.shape[0])
            if self.iters % self.opt.save_latest_freq == 0 and self.rank == 0:
                self.model.save("latest")
        for sched in self.schedulers.values():
            sched.step()

    @torch.no_grad()
    def _val_epoch(self, data_loader):
        """wgan_trainer.py:109-113: critic outputs on validation images (the reference leaves netD in its current mode)."""
        for batch_data in data_loader:
            batch_data = batch_data.to(self.opt.device, non_blocking=True)
            with Fn.weight_cache():
                val_logits = self.model.netD(batch_data)
            self.dis_outputs["val"] += val_logits.flatten().tolist()

    def _make_writer(self):
        if self.rank != 0 or getattr(self.opt, "log_dir", None) is None:
            return None
        try:
            from torch.utils.tensorboard import SummaryWriter
        except Exception:
            return None
        return SummaryWriter(self.opt.log_dir / self.opt.name)

    def _write_tf_log(self, writer, epoch):
        """wgan_trainer.py:54-67."""
        for key, value in self.losses.items():
            writer.add_scalar(f"Loss/{key}", sum(value) / len(value), epoch)
        writer.add_scalars("D(x)", {k: sum(v) / len(v) for k, v in self.dis_outputs.items()}, epoch)
        from torchvision.utils import make_grid
        nrow = int(math.sqrt(self.opt.num_display_images))
        writer.add_image("Generated Image", make_grid(self._generate_image(), nrow=nrow), epoch)

    # ------------------------------------------------------------------------------------------------ the hot path
    def _g_body(self, noise):
        for p in self._d_params:  # the reference computes the critic's gradients here and discards them (D zero_grad precedes their use)
            p.requires_grad_(False)
        try:
            with Fn.weight_cache():
                self.optimizers["G"].zero_grad()
                fake_data = self.model.netG(noise)
                fake_logits = self.model.netD(fake_data)
                g_loss = -fake_logits.mean()
                self._backward(g_loss)
        finally:
            for p in self._d_params:
                p.requires_grad_(True)
        self._reduce_and_step("G")
        return g_loss.detach().reshape(1)

    def _d_body(self, real_data, noise):
        self.model.weight_clipping()
        with Fn.weight_cache():
            self.optimizers["D"].zero_grad()
            with torch.no_grad():
                fake_data = self.model.netG(noise)
            fake_logits = self.model.netD(fake_data.detach_())
            real_logits = self.model.netD(real_data)
            d_loss = -(real_logits.mean() - fake_logits.mean())
            self._backward(d_loss)
        self._reduce_and_step("D")
        return torch.cat([d_loss.detach().reshape(1), real_logits.detach().flatten(), fake_logits.detach().flatten()])

    def _train_generator_once(self, batch_size):
        """wgan_trainer.py:115-122.  ``batch_size`` may also be a noise tensor (netG accepts both, generator.py:41-47)."""
        noise = batch_size if isinstance(batch_size, torch.Tensor) else torch.rand(batch_size, self.opt.noise_dim, 1, 1)  # generator.py:46
        self.losses["gan_G"].append(self._graphed["G"](noise).tolist()[0])

    def _train_discriminator_once(self, real_data, noise=None):
        """wgan_trainer.py:124-138."""
        b = real_data.shape[0]
        if noise is None:
            noise = torch.rand(b, self.opt.noise_dim, 1, 1)  # generator.py:46, drawn by netG(batch_size) in the reference
        vals = self._graphed["D"](real_data, noise).tolist()  # one D2H sync
        self.losses["gan_D"].append(vals[0])
        self.dis_outputs["real"] += vals[1:1 + b]
        self.dis_outputs["fake"] += vals[1 + b:]
