"""DefectGanModel (models/defectgan_model.py:19-104,173-314,385-428) on the sm_100a kernels: the loss graphs of the
'generator' / 'discriminator' / 'inference' / 'inference_classifier' modes; SPADE, SEAN and AdaIN style variants.

Same constructor, ``__call__(mode, data, labels, df_data=None, img_only=False, mask=None)`` contract, return values
and network train/eval switching as the reference; the networks are the mirrored modules of ``models.networks`` whose
forward/backward run libdei2i.so, the losses are ``BaseModel._cal_loss`` (fused reduction kernels).  Packed BF16
weights are built once per loss computation (``Fn.weight_cache``): the 4 G and 2 D passes of a G-step share them.

``opt.diff_aug`` (DiffAugment policies, utils/diffaug.py) runs on the fused augmentation kernels; the MAE pre-training modes
('mae_generator' / 'mae_discriminator' / 'mae_inference', defectgan_model.py:106-171) run the same networks behind the fused
mask-token kernel.
Not on this path: the image-grid helpers and SEAN's `style_distill` (fails in the reference itself, see normalization.SEAN) -
they raise NotImplementedError instead of silently falling back to PyTorch.
"""
import random

import torch

from .. import functional as Fn
from ..utils.diffaug import DiffAugment
from ..utils.util import draw_shifted_mask
from .base_model import BaseModel
from .networks.architecture import MaskToken
from .networks.discriminator import DefectGanDiscriminator
from .networks.extractor import StyleExtractor
from .networks.generator import DefectGanGenerator


def _mean(losses):
    return torch.stack(list(losses)).mean()


class DefectGanModel(BaseModel):
    def __init__(self, opt):
        super().__init__(opt)
        image_size = opt.image_size
        assert image_size & (image_size - 1) == 0, "Image size must be a power of 2"
        if opt.style_norm_block_type not in ("spade", "sean", "adain"):
            raise ValueError(f"|style_norm_block_type {opt.style_norm_block_type}| is invalid")
        self.netG = DefectGanGenerator(opt).to(opt.device, non_blocking=True)
        self.netD = DefectGanDiscriminator(opt).to(opt.device, non_blocking=True)
        if getattr(opt, "is_train", False) or hasattr(opt, "clf_loss_type"):
            assert opt.clf_loss_type is not None, "clf_loss_type should be initialized in dataset"
            self.clf_loss_type = opt.clf_loss_type
        if hasattr(opt, "mask_token_type"):  # learnable mask token of the MAE pre-training path (defectgan_model.py:30-32)
            self.mask_token = MaskToken(opt).to(opt.device, non_blocking=True)
        # style embedding (defectgan_model.py:34-47)
        if opt.style_norm_block_type == "sean":
            if opt.sean_alpha is not None:
                self.netG.set_sean_alpha(opt.sean_alpha)
            if opt.sean_alpha != 0 and not (opt.phase == "test" and opt.use_running_stats):
                assert opt.embed_path is not None, "embed_path should be initialized if style_norm_block_type is sean and sean_alpha is not 0"
                self.embeddings = torch.load(opt.embed_path)
                for label, embeds in self.embeddings.items():
                    self.embeddings[label] = [e.to(opt.device, non_blocking=True) for e in embeds]
        elif opt.style_norm_block_type == "adain":
            self.netE = StyleExtractor(opt).to(opt.device, non_blocking=True)

    # ------------------------------------------------------------------------------------------------ dispatch
    def __call__(self, mode, data, labels, df_data=None, img_only=False, mask=None):
        dev = self.opt.device
        data, labels = data.to(dev, non_blocking=True), labels.to(dev, non_blocking=True)
        if df_data is not None:
            df_data = df_data.to(dev, non_blocking=True)
        if mode.startswith("mae"):  # defectgan_model.py:55-80
            if mode == "mae_generator":
                self.netD.eval()
                self.netG.train()
                with Fn.weight_cache():
                    return self._compute_mae_generator_loss(data, labels)
            if mode == "mae_discriminator":
                self.netD.train()
                self.netG.eval()
                with Fn.weight_cache():
                    return self._compute_mae_discriminator_loss(data, labels)
            if mode == "mae_inference":
                self.netD.eval()
                self.netG.eval()
                with torch.no_grad(), Fn.weight_cache():
                    if self.opt.split_training:
                        rec_loss, gan_loss, _ = self._compute_mae_inference_loss(data, labels)
                        _, clf_loss = self._compute_mae_discriminator_loss(data, labels)
                        return rec_loss, gan_loss, clf_loss
                    return self._compute_mae_inference_loss(data, labels)
            if mode == "mae_generate_grid":
                raise NotImplementedError("mae_generate_grid (TensorBoard image grids, host side) is out of the hot-path scope")
            raise ValueError(f"|mode {mode}| is invalid")
        if mode == "generator":
            self.netD.eval()
            self.netG.train()
            with Fn.weight_cache():
                return self._compute_generator_loss(data, labels, df_data)
        if mode == "discriminator":
            self.netD.train()
            self.netG.eval()
            with Fn.weight_cache():
                return self._compute_discriminator_loss(data, labels, df_data)
        if mode == "inference":
            self.netD.eval()
            self.netG.eval()
            return self._inference(data, labels)
        if mode == "inference_classifier":
            self.netD.eval()
            self.netG.eval()
            return self._compute_clf_loss(data, labels)
        if mode == "generate_grid":
            raise NotImplementedError("generate_grid (TensorBoard image grids, host side) is out of the hot-path scope")
        raise ValueError(f"|mode {mode}| is invalid")

    # ------------------------------------------------------------------------------------------------ loss graphs
    def _compute_generator_loss(self, bg_data, df_labels, df_data):
        """defectgan_model.py:173-249: normal->defect->normal and defect->normal->defect cycles (4 G passes, train mode),
        2 D passes (eval mode), adversarial + classification + reconstruction + spatial-distribution cycle / constraint."""
        nm_labels, nm_feat, df_labels, df_feat = self._get_label_and_style_feat(bg_data, df_labels, df_data)
        track = self.opt.style_norm_block_type == "sean" and self.opt.use_running_stats
        if track:
            self.netG.track_running_stats = True
        fake_defects, df_prob = self.netG(bg_data, df_labels, df_feat)
        recover_normals, rec_df_prob = self.netG(fake_defects, nm_labels, nm_feat)
        fake_normals, nm_prob = self.netG(df_data, nm_labels, nm_feat)
        recover_defects, rec_nm_prob = self.netG(fake_normals, df_labels, df_feat)
        if track:
            self.netG.track_running_stats = False

        policy = getattr(self.opt, "diff_aug", "")  # defectgan_model.py:200-203: the reconstruction terms use the un-augmented images
        fd_src, fd_cls = self.netD(DiffAugment(fake_defects, policy))
        fn_src, fn_cls = self.netD(DiffAugment(fake_normals, policy))

        gan = _mean([self._cal_loss(fd_src, 1.0, "bce"), self._cal_loss(fn_src, 1.0, "bce")])
        clf = _mean([self._cal_loss(fd_cls, df_labels.view_as(fd_cls), self.clf_loss_type),
                     self._cal_loss(fn_cls, nm_labels.view_as(fn_cls), self.clf_loss_type)])
        rec = _mean([self._cal_loss(recover_defects, df_data, "l1"), self._cal_loss(recover_normals, bg_data, "l1")])
        if self.opt.cycle_gan:
            zero = torch.zeros([], device=rec.device)
            return gan, clf, rec, zero, zero.clone()
        cyc = _mean([self._cal_loss(df_prob, rec_df_prob, "l1"), self._cal_loss(nm_prob, rec_nm_prob, "l1")])
        con = _mean([self._cal_loss(p, 0.0, "l1") for p in (df_prob, nm_prob, rec_df_prob, rec_nm_prob)])
        return gan, clf, rec, cyc, con

    def _compute_discriminator_loss(self, bg_data, df_labels, df_data):
        """defectgan_model.py:251-292: 2 G passes without grad (eval mode), 4 D passes (train mode: one spectral-norm
        power iteration each), 4 adversarial + 2 classification terms."""
        nm_labels, nm_feat, df_labels, df_feat = self._get_label_and_style_feat(bg_data, df_labels, df_data)
        with torch.no_grad():
            fake_defects, _ = self.netG(bg_data, df_labels, df_feat)
            fake_normals, _ = self.netG(df_data, nm_labels, nm_feat)
        policy = getattr(self.opt, "diff_aug", "")  # defectgan_model.py:266-271: fakes first, then the real defect / background batches
        fake_defects, fake_normals = DiffAugment(fake_defects, policy), DiffAugment(fake_normals, policy)
        df_data, bg_data = DiffAugment(df_data, policy), DiffAugment(bg_data, policy)
        fd_src, _ = self.netD(fake_defects.detach_())
        fn_src, _ = self.netD(fake_normals.detach_())
        rd_src, rd_cls = self.netD(df_data)
        rn_src, rn_cls = self.netD(bg_data)
        gan = _mean([self._cal_loss(fd_src, 0.0, "bce"), self._cal_loss(fn_src, 0.0, "bce"),
                     self._cal_loss(rd_src, 1.0, "bce"), self._cal_loss(rn_src, 1.0, "bce")])
        clf = _mean([self._cal_loss(rd_cls, df_labels.view_as(rd_cls), self.clf_loss_type),
                     self._cal_loss(rn_cls, nm_labels.view_as(rn_cls), self.clf_loss_type)])
        return gan, clf

    # ------------------------------------------------------------------------------------------------ MAE pre-training
    def _repair_mask(self, imgs, labels, mask=None):
        """defectgan_model.py:361-383 (spade branch): shifted patch mask -> mask token -> generator.  The mask is drawn with the
        reference's host RNG calls (utils.util.draw_shifted_mask) and blended on the device in one pass (MaskToken.blend)."""
        if mask is None:
            coarse, hs, ws = draw_shifted_mask(imgs.size(), self.opt.patch_size, self.opt.mask_ratio)
            masked_imgs, masks = self.mask_token.blend(imgs, coarse, hs, ws, self.opt.patch_size)
        else:
            masks = mask.to(self.opt.device, non_blocking=True)
            masked_imgs = self.mask_token(imgs, masks)
        kind = self.opt.style_norm_block_type
        style_feat = self._get_style_embeds(labels) if kind == "sean" else self.netE(imgs, labels) if kind == "adain" else None
        predicted_imgs, _ = self.netG(masked_imgs, self._expand_seg(labels), style_feat)
        return predicted_imgs, masks

    def _mae_adv_terms(self, predicted_imgs, labels):
        fake_src, fake_cls = self.netD(predicted_imgs)
        return self._cal_loss(fake_src, 1.0, "bce"), self._cal_loss(fake_cls, labels, self.clf_loss_type)

    def _compute_mae_generator_loss(self, imgs, labels):
        """defectgan_model.py:106-132: L1 reconstruction of the full image from the masked one (+ adversarial / classification terms
        through the discriminator unless --split_training).  Returns (rec, gan, clf)."""
        predicted_imgs, _ = self._repair_mask(imgs, labels)
        rec_loss = self._cal_loss(predicted_imgs, imgs, "l1")
        if self.opt.split_training:
            zero = torch.zeros([], device=rec_loss.device)
            return rec_loss, zero, zero.clone()
        gan_loss, clf_loss = self._mae_adv_terms(predicted_imgs, labels)
        return rec_loss, gan_loss, clf_loss

    @torch.no_grad()
    def _compute_mae_inference_loss(self, imgs, labels):
        """defectgan_model.py:134-150."""
        predicted_imgs, _ = self._repair_mask(imgs, labels)
        rec_loss = self._cal_loss(predicted_imgs, imgs, "l1")
        if self.opt.split_training:
            zero = torch.zeros([], device=rec_loss.device)
            return rec_loss, zero, zero.clone()
        gan_loss, clf_loss = self._mae_adv_terms(predicted_imgs, labels)
        return rec_loss, gan_loss, clf_loss

    def _compute_mae_discriminator_loss(self, imgs, labels):
        """defectgan_model.py:152-171: real images -> classification (+ real / repaired adversarial terms unless --split_training).
        Returns (gan, clf)."""
        real_src, real_cls = self.netD(imgs)
        clf_loss = self._cal_loss(real_cls, labels, self.clf_loss_type)
        if self.opt.split_training:
            return torch.zeros([], device=clf_loss.device), clf_loss
        with torch.no_grad():
            predicted_imgs, _ = self._repair_mask(imgs, labels)
        fake_src, _ = self.netD(predicted_imgs.detach_())
        gan = _mean([self._cal_loss(fake_src, 0.0, "bce"), self._cal_loss(real_src, 1.0, "bce")])
        return gan, clf_loss

    def _compute_clf_loss(self, imgs, labels):
        """defectgan_model.py:294-300."""
        _, logits = self.netD(imgs)
        return logits, self._cal_loss(logits, labels, self.clf_loss_type)

    @torch.no_grad()
    def _inference(self, data, labels):
        """``_generate_fake`` under no_grad as a CUDA graph per input-shape signature (configs[4]: at batch <= 16 the ~150
        launches of a generator pass are host-bound).  Call 1 of a signature runs eagerly, call 2 captures, later calls copy
        the inputs into the static buffers and replay; the weight packing and the eval-mode spectral-norm sigma are part of the
        graph and read the parameters in place, so training between two calls is seen.  DEI2I_CUDA_GRAPH=0 disables."""
        import os
        graphs = self.__dict__.setdefault("_inf_graphs", {})
        use = (os.environ.get("DEI2I_CUDA_GRAPH", "1") != "0" and data.is_cuda and not torch.is_grad_enabled()
               and not torch.cuda.is_current_stream_capturing())
        if not use:
            with Fn.weight_cache():
                return self._generate_fake(data, labels)
        key = (tuple(data.shape), tuple(labels.shape), data.dtype, labels.dtype)
        st = graphs.get(key)
        if st is None:
            graphs[key] = {}
            with Fn.weight_cache():
                return self._generate_fake(data, labels)
        if "graph" not in st:
            st["x"], st["l"] = data.clone(), labels.clone()
            torch.cuda.synchronize()
            g = torch.cuda.CUDAGraph()
            with torch.cuda.graph(g), Fn.weight_cache():
                st["out"] = self._generate_fake(st["x"], st["l"])
            st["graph"] = g
        else:
            st["x"].copy_(data, non_blocking=True)
            st["l"].copy_(labels, non_blocking=True)
        st["graph"].replay()
        return tuple(o.clone() for o in st["out"])

    def _generate_fake(self, data, labels):
        """defectgan_model.py:302-314: labels [B,label_nc] or (spade) a spatial map [B,label_nc,h,w]."""
        kind = self.opt.style_norm_block_type
        style_feat = self._get_style_embeds(labels) if kind == "sean" else self.netE(data, labels) if kind == "adain" else None
        return self.netG(data, self._expand_seg(labels), style_feat)

    # ------------------------------------------------------------------------------------------------ helpers
    def _cal_loss(self, logits, targets, loss_type):
        """A python scalar target means a constant target map (the reference materialises ones_like / zeros_like)."""
        if isinstance(targets, (int, float)) and loss_type == "l1":
            targets = None if targets == 0 else torch.full_like(logits, float(targets))
        return super()._cal_loss(logits, targets, loss_type)

    def _expand_seg(self, labels):
        if labels.dim() == 2:
            return labels.reshape(labels.size(0), labels.size(1), 1, 1)
        if labels.dim() == 4:
            return labels
        raise ValueError(f"|labels dim {labels.dim()}| is invalid")

    def _labels(self, df_labels):
        """defectgan_model.py:413-416: normal label = one-hot class 0; both as [B,label_nc,1,1] maps (the generator reads the rows
        back for the vector-conditioned SEAN / AdaIN norms)."""
        df_labels = df_labels.float()
        nm_labels = torch.zeros_like(df_labels)
        nm_labels[:, 0] = 1
        return self._expand_seg(nm_labels), self._expand_seg(df_labels)

    def _get_style_embeds(self, labels):
        """defectgan_model.py:394-411: `num_embeds` stored embeddings per sample, drawn (python `random.choices`, as the reference)
        from the list kept for the sample's label combination; zeros when that list is empty."""
        if self.opt.sean_alpha == 0:
            return None
        if self.opt.use_running_stats and self.netG.inference_running_stats:
            return torch.randn(labels.size(0), self.opt.hidden_nc).to(self.netG.device)
        num_embeds = self.opt.num_embeds
        rows = []
        for label in labels.reshape(labels.size(0), -1).int().tolist():
            pool = self.embeddings[tuple(label)]
            if not pool:
                rows.append(torch.zeros(num_embeds, self.opt.embed_nc, device=self.netG.device))
            else:
                rows.append(torch.stack(random.choices(pool, k=num_embeds)))
        return torch.stack(rows)

    def _get_label_and_style_feat(self, bg_data, df_labels, df_data):
        """defectgan_model.py:413-428."""
        kind = self.opt.style_norm_block_type
        nm_rows = torch.zeros_like(df_labels, dtype=torch.float)
        nm_rows[:, 0] = 1
        nm_feat = df_feat = None
        if kind == "sean":
            nm_feat = self._get_style_embeds(nm_rows)
            df_feat = self._get_style_embeds(df_labels)
        elif kind == "adain":
            nm_feat = self.netE(bg_data, nm_rows)
            df_feat = self.netE(df_data, df_labels)
        nm_labels, df_labels = self._labels(df_labels)
        if kind == "spade":  # both label tensors of the loss graph go through each SPADE module's class-map convolutions together
            Fn.announce_label_sets(df_labels, nm_labels)
        return nm_labels, nm_feat, df_labels, df_feat
