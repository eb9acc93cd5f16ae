"""DefectGanGenerator (models/networks/generator.py:52-328) on the sm_100a kernels; SPADE / SEAN / AdaIN variants, skip_conn=False."""
import math

import torch
from torch import nn

from ... import functional as Fn
from ...ops import ConvSpec
from .architecture import (ConvBlock, DeConvBlock, NormConvBlock, NormResBlock, ResBlock, _consumer_spec, _edge, _fusable_pad, _spec_of,
                           next_pad, same_frame)
from .base_network import BaseNetwork
from .normalization import SEAN


class WGanGenerator(BaseNetwork):
    """models/networks/generator.py:11-49: Upsample, [Upsample + 4x4 'same' conv + BN + ReLU] x (num_layers + 1), Upsample,
    4x4 'same' conv, tanh.  Same ``noise_to_image`` Sequential (state-dict keys, init order) as the reference."""

    def __init__(self, num_layers, ngf=64, noise_dim=100):
        super().__init__()
        self.noise_dim = noise_dim
        model = [nn.Upsample(scale_factor=2)]
        dim = ngf * (2 ** num_layers)
        model += [DeConvBlock(noise_dim, dim, kernel_size=(4, 4), padding="same", norm_layer=nn.BatchNorm2d, act_layer="relu")]
        for _ in range(num_layers):
            model += [DeConvBlock(dim, dim // 2, kernel_size=(4, 4), stride=(1, 1), padding="same", norm_layer=nn.BatchNorm2d,
                                  act_layer="relu")]
            dim //= 2
        model += [nn.Upsample(scale_factor=2), nn.Conv2d(dim, 3, kernel_size=(4, 4), stride=(1, 1), padding="same", bias=False),
                  nn.Tanh()]
        self.noise_to_image = nn.Sequential(*model)
        self._out_spec = _spec_of(self.noise_to_image[-2])
        self._out_frame = same_frame(self.noise_to_image[-2])

    def forward(self, x):
        assert isinstance(x, (torch.Tensor, int)), "x must be batch_size (Int) or noise (Torch.Tensor)"
        noise = x if isinstance(x, torch.Tensor) else torch.rand(x, self.noise_dim, 1, 1)  # uniform, generator.py:46
        noise = noise.to(self.device)
        mods = list(self.noise_to_image)
        feat = Fn.upsample_pad(Fn.to_nhwc(noise.float()), 2)
        for blk in mods[1:-3]:
            feat = blk(feat)
        conv = mods[-2]
        feat = Fn.upsample_pad(feat, 2, self._out_frame or (0, 0, 0, 0))
        y = Fn.conv(feat, conv, self._out_spec, conv.training, act="tanh", out_fp32=True)
        return _edge(y)


class DefectGanGenerator(BaseNetwork):
    def __init__(self, opt):
        super().__init__()
        assert (opt.num_res & 1) == 0, "num_res must be even"
        if getattr(opt, "skip_conn", False):
            raise NotImplementedError("skip_conn=True is broken in the reference (architecture.py:451 TODO) and not mirrored")
        self.opt = opt
        self.cycle_gan = opt.cycle_gan
        self.label_nc = opt.label_nc
        self.skip_conn = False
        sn = opt.use_spectral
        dim = opt.ngf
        self.stem = ConvBlock(opt.input_nc, dim, kernel_size=(7, 7), padding="same", padding_mode="reflect",
                              norm_layer=nn.BatchNorm2d, act_layer="leaky_relu", use_spectral=sn)
        enc, enc_res, dec_res, dec = [], [], [], []
        for _ in range(opt.num_scales):
            enc.append(ConvBlock(dim, dim * 2, kernel_size=(4, 4), stride=(2, 2), padding=1, padding_mode="reflect",
                                 norm_layer=nn.BatchNorm2d, act_layer="leaky_relu", use_spectral=sn))
            dim *= 2
        for _ in range(opt.num_res // 2):
            enc_res.append(ResBlock(dim, dim, kernel_size=(3, 3), stride=(1, 1), padding="same", padding_mode="reflect",
                                    norm_layer=nn.BatchNorm2d, act_layer="leaky_relu", use_spectral=sn))
        style = dict(style_distill=opt.style_distill, embed_nc=opt.embed_nc, kernel_size=(3, 3), stride=(1, 1),
                     padding="same", padding_mode="reflect", norm_layer=nn.InstanceNorm2d, act_layer="relu",
                     use_spectral=sn, add_noise=opt.add_noise)
        for _ in range(opt.num_res // 2, opt.num_res):
            dec_res.append(NormResBlock(opt.style_norm_block_type, opt.hidden_nc, opt.label_nc, dim, dim, up_scale=False, **style))
        head = dict(kernel_size=(3, 3), padding="same", padding_mode="reflect", up_scale=False, norm_layer=None,
                    use_spectral=False, add_noise=False)
        for _ in range(opt.num_scales):
            dec.append(NormConvBlock(opt.style_norm_block_type, opt.hidden_nc, opt.label_nc, dim, dim // 2, up_scale=True, **style))
            dim //= 2
            # The reference (re)builds the containers and both heads inside this loop (generator.py:206-241), so the
            # default-initialiser RNG draws of the discarded heads happen there too; mirrored so that the same seed
            # gives the same parameters.  Registration order: enc_blk, enc_res_blk, dec_res_blk, dec_blk, heads.
            self.enc_blk = nn.Sequential(*enc)
            self.enc_res_blk = nn.Sequential(*enc_res)
            self.dec_res_blk = nn.Sequential(*dec_res)
            self.dec_blk = nn.Sequential(*dec)
            self.foreground_head = DeConvBlock(dim, 3, act_layer="tanh", **head)
            self.distribution_head = DeConvBlock(dim, 1, act_layer="sigmoid", **head)
        self._head_spec = ConvSpec(dim, 4, 3, 1, 1, "reflect")

    def forward(self, x, labels, style_feat=None):
        """(output, spatial_prob) like generator.py:243-275; x: [B,3,H,W] FP32 in [-1,1]; labels: [B,label_nc,1,1] or a
        spatial map [B,label_nc,h,w].  Both heads are one N=4 convolution whose FP32 logits feed the fused
        tanh / sigmoid / composition kernel (generator.py:268-270)."""
        assert isinstance(x, torch.Tensor), "x must be Original Images: Torch.Tensor"
        if labels.dim() == 2:
            labels = labels.reshape(labels.size(0), labels.size(1), 1, 1)
        seg = Fn.to_nhwc(labels.float())
        seg.key = Fn.label_key(labels)
        Fn.prefetch_spectral_norm(self, self.training)
        Fn.prefetch_packs(self)
        enc = list(self.enc_blk)
        feat = self.stem(x, carrier_out=True, out_pad=next_pad([self.stem] + enc, 0))
        for i, blk in enumerate(enc):
            feat = blk(feat, labels, out_pad=next_pad(enc, i)) if isinstance(blk, ConvBlock) else blk(feat, labels)
        for blk in self.enc_res_blk:
            feat = blk(feat, labels)
        # every decoder block but the last feeds a SPADE (its per-image statistics come out of the convolution epilogue); the last
        # one feeds the 3x3 reflect heads (their padding comes out of it)
        dec = list(self.dec_blk)
        for blk in self.dec_res_blk:
            feat = blk(feat, seg, style_feat, stats_out=True)
        for i, blk in enumerate(dec):
            last = i + 1 == len(dec)
            feat = blk(feat, seg, style_feat, stats_out=not last, out_pad=_fusable_pad(self._head_spec) if last else 0)
        # NOTE generator.py:266-267 (host-synchronising isnan().any() + nan_to_num_) is not reproduced: it is the
        # identity on finite activations and would cost a device->host sync per forward.
        hd = Fn.conv_cat(feat, self.foreground_head._conv, self.distribution_head._conv, _consumer_spec(self._head_spec, feat), out_fp32=True)
        return Fn.compose(hd.t, x, self.cycle_gan)

    # ---- SEAN bookkeeping (generator.py:276-328) ----
    def _seans(self):
        return [m for m in self.modules() if isinstance(m, SEAN)]

    def update_per_epoch(self, epoch):
        super().update_per_epoch(epoch)
        alpha = (1 + math.cos(math.pi * epoch / self.opt.num_epochs)) / 2
        if self.opt.style_norm_block_type == "sean":
            if self.opt.sean_alpha is None:
                self.set_sean_alpha(alpha)
            if self.opt.use_running_stats:
                self.update_stats()

    def set_sean_alpha(self, alpha):
        for m in self._seans():
            m.set_alpha(alpha)

    def enable_sean_distill_loss(self, enable_distill_loss):
        for m in self._seans():
            m.distill_loss = enable_distill_loss

    def get_sean_distill_loss(self):
        return {"latent": [], "embed": []}  # style_distill is not mirrored (normalization.SEAN docstring)

    def update_stats(self):
        for m in self._seans():
            m.update_stats()

    @property
    def track_running_stats(self):
        for m in self._seans():
            return m.track_running_stats

    @track_running_stats.setter
    def track_running_stats(self, value):
        for m in self._seans():
            m.track_running_stats = value

    @property
    def inference_running_stats(self):
        for m in self._seans():
            return m.inference_running_stats

    @inference_running_stats.setter
    def inference_running_stats(self, value):
        for m in self._seans():
            m.inference_running_stats = value
