"""MAETrainer (trainers/mae_trainer.py:13-158): masked-autoencoder pre-training of the DefectGAN generator / discriminator - the
first stage of the README's workflow (defectGAN/README.md:6-18).

Same step bodies as the reference: D-step = classification of real images (+ real / repaired adversarial terms), G-step =
L1 reconstruction of the full image from its masked version (+ adversarial / classification through D), loss weights
[rec, cls_d, cls_g] = ``opt.loss_weight``, the mask token's parameters added to the generator's optimizer as a second group.
Differences, invisible to the results: FlatAdam (AdamW semantics for ``--optimizer adamw``, the pre-training default) on the fused
kernel; the reference's ``torch.cuda.amp.GradScaler`` (mae_trainer.py:134-137,150-152) guards FP16 autocast - this path computes in
BF16 with FP32 accumulation, which needs no loss scaling, so the scaler's unscale / inf-check / update are no-ops and omitted.
The TensorBoard image grids ('mae_generate_grid') and the FID validation are host-side and out of the hot-path scope."""
from collections import defaultdict

import numpy as np
import torch

from .base_trainer import BaseTrainer


class MAETrainer(BaseTrainer):
    def __init__(self, opt, data_types=()):
        super().__init__(opt)
        self.loss_weights = {"rec": opt.loss_weight[0], "clf_D": opt.loss_weight[1], "clf_G": opt.loss_weight[2]}
        self.loss_types = ["rec", "gan", "clf"]
        self._init_losses()
        self._has_e = "E" in self.optimizers  # AdaIN variant: the StyleExtractor is stepped with the generator (mae_trainer.py:123-139)
        self.data_types = data_types
        # add mask_token's param (mae_trainer.py:28)
        self.optimizers["G"].add_param_group({"params": self.model.mask_token.parameters()})
        self._d_params = list(self.model.netD.parameters())

    def _init_lr(self, opt):
        assert len(opt.lr) in (1, 2), f"length of lr must be 1 or 2, not {len(opt.lr)}"
        self.lr = {"D": opt.lr[0], "G": opt.lr[1]} if len(opt.lr) == 2 else opt.lr[0]
        if opt.style_norm_block_type == "adain" and isinstance(self.lr, dict):  # mae_trainer.py:34-35 (which fails on a scalar lr)
            self.lr["E"] = opt.lr[1]

    def _init_losses(self):
        self.losses = {t: defaultdict(list) for t in self.loss_types}

    # ------------------------------------------------------------------------------------------------ the hot path
    def _train_generator_once(self, data, labels):
        """mae_trainer.py:121-146."""
        self.optimizers["G"].zero_grad()
        if self._has_e:
            self.optimizers["E"].zero_grad()
        for p in self._d_params:  # the reference computes D's gradients here and discards them at D's next zero_grad
            p.requires_grad_(False)
        try:
            rec_loss, gan_loss, clf_loss = self.model("mae_generator", data, labels)
            g_loss = gan_loss + rec_loss * self.loss_weights["rec"] + clf_loss * self.loss_weights["clf_G"]
            self._backward(g_loss)
        finally:
            for p in self._d_params:
                p.requires_grad_(True)
        self._reduce_and_step("G")
        if self._has_e:
            self._reduce_and_step("E")
        vals = torch.stack([rec_loss.detach(), gan_loss.detach(), clf_loss.detach()]).tolist()
        self.losses["rec"]["train"].append(vals[0])
        self.losses["gan"]["G"].append(vals[1])
        self.losses["clf"]["G"].append(vals[2])

    def _train_discriminator_once(self, data, labels):
        """mae_trainer.py:148-158."""
        self.optimizers["D"].zero_grad()
        gan_loss, clf_loss = self.model("mae_discriminator", data, labels)
        d_loss = gan_loss + clf_loss * self.loss_weights["clf_D"]
        self._backward(d_loss)
        self._reduce_and_step("D")
        vals = torch.stack([gan_loss.detach(), clf_loss.detach()]).tolist()
        self.losses["gan"]["D"].append(vals[0])
        self.losses["clf"]["D"].append(vals[1])

    # ------------------------------------------------------------------------------------------------ epoch loop
    def train(self, train_loaders, val_loaders=None):
        """mae_trainer.py:72-87; epochs run from first_epoch to num_epochs inclusive."""
        writer = self._make_writer()
        for epoch in range(self.first_epoch, self.opt.num_epochs + 1):
            self._init_losses()
            self._train_epoch(train_loaders, epoch)
            if self.opt.phase == "val" and val_loaders is not None:
                self._val_epoch(val_loaders, epoch)
            if writer is not None:
                self._write_tf_log(writer, epoch)
            if epoch % self.opt.save_ckpt_freq == 0 and self.rank == 0:
                self.model.save(epoch)
            self._update_per_epoch(epoch)
        if writer is not None:
            writer.close()

    def _train_epoch(self, data_loaders, epoch):
        """mae_trainer.py:89-112 (progress bar omitted)."""
        for bg_data, bg_labels, _ in data_loaders["fusion"]:
            self.iters += 1
            self._train_discriminator_once(bg_data, bg_labels)
            if self.iters % self.opt.num_critics == 0:
                self._train_generator_once(bg_data, bg_labels)
            if self.iters % self.opt.save_latest_freq == 0 and self.rank == 0:
                self.model.save("latest")
                np.savetxt(self.iter_record_path, (epoch, self.iters), fmt="%i", delimiter=",")

    @torch.no_grad()
    def _val_epoch(self, data_loader, epoch):
        """mae_trainer.py:114-119."""
        for data, labels, _ in data_loader["fusion"]:
            rec_loss, gan_loss, clf_loss = self.model("mae_inference", data, labels)
            self.losses["rec"]["val"].append(rec_loss.item())
            self.losses["gan"]["val"].append(gan_loss.item())
            self.losses["clf"]["val"].append(clf_loss.item())

    def _make_writer(self):
        if self.rank != 0 or getattr(self.opt, "log_dir", None) is None:
            return None
        try:
            from torch.utils.tensorboard import SummaryWriter
        except Exception:
            return None
        return SummaryWriter(self.opt.log_dir / self.opt.name)

    def _write_tf_log(self, writer, epoch):
        """mae_trainer.py:42-70 without the image grids."""
        for t in self.loss_types:
            writer.add_scalars(f"Losses/{t}", {k: sum(v) / (len(v) + 1e-12) for k, v in self.losses[t].items()}, epoch)
        for name, sched in self.schedulers.items():
            writer.add_scalar(f"Lr/{name}", sched.get_last_lr()[0], epoch)
        if self.opt.mask_token_type in ("scalar", "vector", "position", "full"):
            tok = self.model.mask_token.mask_token.detach().clone().squeeze(0).cpu()
            tok = tok.expand(self.opt.input_nc, self.opt.image_size, self.opt.image_size)
            writer.add_image("Images/mask_token", torch.clamp(tok, -1, 1).add(1).div(2), epoch)

    def _update_per_epoch(self, epoch=None):
        super()._update_per_epoch(epoch)
        self.model.update_per_epoch(epoch)
