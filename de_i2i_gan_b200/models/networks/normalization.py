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

    def _gb5(self, cls5):
        """mlp_shared -> ReLU -> (mlp_gamma | mlp_beta as one N = 2C convolution) on [n, 5, 5, label_nc] border-class maps."""
        sconv = self.mlp_shared[0]
        actv = Fn.conv(cls5, sconv, self._shared_spec, self.training, bias=sconv.bias, act="relu")
        return Fn.conv_cat(actv, self.mlp_gamma, self.mlp_beta, self._gb_spec)

    def _gb5_for_label_sets(self, seg, n, ck):
        """All label tensors announced for this scope (Fn.announce_label_sets) in one batched evaluation; each set's slice is cached
        under its own key, the one asked for is returned.  None when `seg` is not one of the announced sets."""
        sets = Fn._cache.get("label_sets")
        if not sets or seg.key not in [Fn.label_key(t) for t in sets] or any(tuple(t.shape) != tuple(sets[0].shape) for t in sets):
            return None
        cat = Fn._cache.get(("cls5cat", n))
        if cat is None:
            maps = [Fn.to_nhwc(t.float()) for t in sets]
            cat = Fn.Nhwc(torch.cat([m.t.expand(n, 5, 5, m.t.shape[3]) for m in maps], dim=0).contiguous(), maps[0].c)
            Fn._cache[("cls5cat", n)] = cat
        gb_all = self._gb5(cat)
        mine = None
        for t, part in zip(sets, torch.split(gb_all.t, n, dim=0)):
            k = ("gb5", id(self), Fn.label_key(t)) + ck[3:]
            Fn._cache[k] = Fn.Nhwc(part, gb_all.c)
            if k == ck:
                mine = Fn._cache[k]
        return mine

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
        if CONSTANT_LABEL_CLASSES and seg.t.shape[1] == 1 and seg.t.shape[2] == 1 and h >= 5 and w >= 5 \
                and self._shared_spec.k == 3 and self._shared_spec.pad == 1:
            # 1x1 label map (every training call, defectgan_model.py:185-205): the nearest-resized map is constant, so
            # mlp_shared -> mlp_gamma|mlp_beta are evaluated on the 5x5 border-class map (exact; dei2i_spade_cls_fwd)
            # gamma|beta depend on the label tensor and this module's weights only: the four generator passes of a G-step use
            # two label tensors (df_labels, nm_labels, nm_labels, df_labels - defectgan_model.py:185-205), so inside one
            # weight_cache() scope (= one loss computation, weights fixed) each is evaluated once and shared; autograd sums
            # the gradients of the passes that used it before the one backward through the two small convolutions
            ck = None
            if Fn._cache is not None and seg.key is not None:
                ck = ("gb5", id(self), seg.key, n, self.training, torch.is_grad_enabled())
            gb5 = Fn._cache.get(ck) if ck is not None else None
            if gb5 is None and ck is not None:
                gb5 = self._gb5_for_label_sets(seg, n, ck)
            if gb5 is None:
                if seg.cls5 is None:
                    seg.cls5 = Fn.Nhwc(seg.t.expand(n, 5, 5, seg.t.shape[3]).contiguous(), seg.c)
                gb5 = self._gb5(seg.cls5)
                if ck is not None:
                    Fn._cache[ck] = gb5
            pad = 0 if plain else out_pad
            out = Fn.spade_modulate_cls(xa, gb5, up=up, relu=relu, eps=self.param_free_norm.eps, out_pad=pad)
            out.pad = pad
            return Fn.from_nhwc(out) if plain else out
        sconv = self.mlp_shared[0]
        actv = Fn.conv(seg, sconv, self._shared_spec, self.training, bias=sconv.bias, act="relu", logical_hw=(h, w))
        gb = Fn.conv_cat(actv, self.mlp_gamma, self.mlp_beta, self._gb_spec)
        out = Fn.spade_modulate(xa, gb, up=up, relu=relu, eps=self.param_free_norm.eps)
        return Fn.from_nhwc(out) if plain else out


def _instance_norm_only(norm):
    if not isinstance(norm, nn.InstanceNorm2d) or norm.affine or norm.track_running_stats:
        raise NotImplementedError("the CUDA path implements the style norms over InstanceNorm2d(affine=False) (the DefectGAN configuration)")


def _modulate(x, gamma, beta, eps, up, relu):
    """relu?((1 + gamma[n,c]) * IN(x) + beta[n,c]) in one pass over x (dei2i_adain_act_fwd), then the 2x nearest up-sampling of the
    consuming NormConvBlock: InstanceNorm statistics are invariant under nearest up-sampling and the modulation is point-wise, so
    up(x) -> norm -> act == norm -> act -> up(x) exactly, and the normalisation reads a quarter of the pixels."""
    plain = not isinstance(x, Fn.Nhwc)
    xa = Fn.to_nhwc(x)
    y = Fn.adain_act(xa, 1 + gamma, beta, slope=0.0 if relu else 1.0, eps=eps)
    if up > 1:
        y = Fn.upsample_pad(y, up)
    return Fn.from_nhwc(y) if plain else y


class AdaIN(nn.Module):
    """normalization.py:40-73, denorm_type='linear': gamma / beta are Linear(hidden_nc -> C) of the style vector."""

    def __init__(self, norm_nc, hidden_nc=None, norm_layer=nn.BatchNorm2d, denorm_type="linear"):
        super().__init__()
        self.denorm_type = denorm_type
        self.param_free_norm = norm_layer(norm_nc, affine=False)
        _instance_norm_only(self.param_free_norm)
        if denorm_type != "linear":
            raise NotImplementedError("AdaIN(denorm_type='stat') is never constructed by the reference (architecture.py:224-225,315-320)")
        assert hidden_nc is not None, "hidden_nc should be set when denorm_type is linear."
        self.hidden_nc = hidden_nc
        self.mlp_gamma = nn.Linear(hidden_nc, norm_nc)
        self.mlp_beta = nn.Linear(hidden_nc, norm_nc)

    def forward(self, x, style_feat, up=1, relu=False):
        n = x.shape[0]
        assert style_feat.size(1) == self.hidden_nc, "The channel of style feature is not equal to hidden_nc."
        s = style_feat.reshape(n, self.hidden_nc).float()
        return _modulate(x, Fn.linear(s, self.mlp_gamma), Fn.linear(s, self.mlp_beta), self.param_free_norm.eps, up, relu)


def label_to_str(label):
    """utils/util.py:178-180."""
    return "-".join(str(i) for i, v in enumerate(label) if v == 1)


class SEAN(nn.Module):
    """normalization.py:76-202.  gamma / beta = Linear(mix) with mix = mean_e(ReLU(Linear(embeds[n, e]))) + ReLU(Linear(labels)) (or the
    label code alone when no embeddings are given); the running per-label statistics (buffers mean_* / std_*, `update_stats`,
    `inference_running_stats`) are kept as the reference keeps them, the per-sample Python loops vectorised.

    `style_distill=True` is not mirrored: with the 3-D embeddings `DefectGanModel._get_style_embeds` always produces, the
    reference's own `calc_kl_with_logits(enc_feat [N,E,H], mix_labels [N,H])` (normalization.py:189-190) raises a shape error."""

    def __init__(self, embed_nc, norm_nc, label_nc, hidden_nc=128, latent_dim=16, norm_layer=nn.BatchNorm2d, style_distill=False):
        super().__init__()
        if style_distill:
            raise NotImplementedError("SEAN(style_distill=True) fails in the reference itself (normalization.py:189-190 on 3-D embeddings)")
        self.style_distill = False
        self.latent_dim = latent_dim
        self.noise_dim = latent_dim - label_nc
        self.label_nc = label_nc
        self.hidden_nc = hidden_nc
        self.alpha = 1.0
        self.param_free_norm = norm_layer(norm_nc, affine=False)
        _instance_norm_only(self.param_free_norm)
        self.mlp_shared = nn.Sequential(nn.Linear(embed_nc, hidden_nc), nn.ReLU(inplace=True))
        self.mlp_gamma = nn.Linear(hidden_nc, norm_nc)
        self.mlp_beta = nn.Linear(hidden_nc, norm_nc)
        self.mlp_latent = nn.Sequential(nn.Linear(label_nc, hidden_nc), nn.ReLU(inplace=True))
        self.create_stats()
        self.inference_running_stats = False
        self.track_running_stats = False
        self.num_embeds_tracked = 10000

    # ---- running statistics of the mixed style code per label combination (normalization.py:106-125) ----
    def create_stats(self):
        combos = torch.cartesian_prod(*([torch.tensor([0, 1])] * self.label_nc)).tolist()
        for l in combos:
            self.register_buffer("mean_" + label_to_str(l), torch.zeros(self.hidden_nc))
            self.register_buffer("std_" + label_to_str(l), torch.zeros(self.hidden_nc))
        self.embeds = {tuple(l): [] for l in combos}

    def update_stats(self):
        eps = 1e-5
        for l, feats in self.embeds.items():
            if feats:
                mean = getattr(self, "mean_" + label_to_str(l))
                std = getattr(self, "std_" + label_to_str(l))
                feat = torch.stack(feats, dim=0)
                # the reference stores them crosswise (`mean[:], std[:] = new_std, new_mean`, normalization.py:124); kept, because
                # checkpoints written by it are read back the same way
                mean[:], std[:] = (feat.var(dim=0) + eps).sqrt(), feat.mean(dim=0)
                self.embeds[l] = feats[-self.num_embeds_tracked:]

    @property
    def distill_loss(self):
        return None

    @distill_loss.setter
    def distill_loss(self, enable):
        pass

    def set_alpha(self, alpha):
        self.alpha = alpha

    def update_alpha(self, epoch, num_epochs):
        pass

    def _label_rows(self, labels):
        return [tuple(int(v) for v in row) for row in labels.detach().cpu().tolist()]

    def forward(self, x, labels, feat=None, up=1, relu=False):
        n = x.shape[0]
        if labels.dim() == 4:
            labels = labels.reshape(n, -1)
        labels = labels.float()
        latent_code = Fn.linear(labels, self.mlp_latent[0], act="relu")
        if feat is None:
            mix = latent_code
        elif self.inference_running_stats:
            rows = self._label_rows(labels)
            mean = torch.stack([getattr(self, "mean_" + label_to_str(l)) for l in rows])
            std = torch.stack([getattr(self, "std_" + label_to_str(l)) for l in rows])
            mix = feat.float() * std * 1.5 + mean
        else:
            f = feat.float()
            enc = Fn.linear(f.reshape(-1, f.shape[-1]), self.mlp_shared[0], act="relu").reshape(*f.shape[:-1], self.hidden_nc)
            mix = enc + latent_code.view(n, 1, -1)
            if mix.dim() == 3:
                mix = mix.mean(dim=1)
            if self.track_running_stats:
                for l, single in zip(self._label_rows(labels), mix.detach().clone()):
                    self.embeds[l].append(single)
            empty = (mix == 0).all(dim=1, keepdim=True)
            mix = torch.where(empty, latent_code, mix)
        return _modulate(x, Fn.linear(mix, self.mlp_gamma), Fn.linear(mix, self.mlp_beta), self.param_free_norm.eps, up, relu)
