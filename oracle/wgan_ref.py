"""TEST INFRASTRUCTURE ONLY - plain-torch FP32 CPU restatement of the WGAN path (BASELINE.json configs[2]).

Networks are re-stated functionally over reference-format ``state_dict``s of ``WGanGenerator`` /
``WGanDiscriminator`` (/root/reference/defectGAN/models/networks/generator.py:11-49, discriminator.py:10-46);
the training step follows trainers/wgan_trainer.py:115-138 and models/wgan_model.py:18-21.  The reference's
``WGanTrainer`` cannot be constructed as checked in, so the 15-line step is restated around the
reference's own networks; pinned in tests/test_oracle_vs_reference.py (live) and tests/golden/wgan_small.pt.
The reference has NO gradient-penalty implementation (only the commented-out ``_cal_dis_grad`` probe,
wgan_trainer.py:30-39): a WGAN-GP variant is "parity unpinned" and is not part of this oracle.
Never imported by the product package.
"""
from __future__ import annotations

from typing import List, Sequence

import torch
import torch.nn.functional as F

from .defectgan_ref import State, Tensor, act, batch_norm, conv2d, trainable_keys


def num_layers_for(image_size: int) -> int:
    """models/wgan_model.py:12-13."""
    assert image_size & (image_size - 1) == 0, "Image size must be a power of 2"
    return image_size.bit_length() - 1 - 3


def same_conv_even(x: Tensor, w: Tensor) -> Tensor:
    """nn.Conv2d(padding='same') with an even kernel (generator.py:18-21,24-29,32-36): torch pads
    total = k - 1 zeros, left/top = total // 2 and the remainder right/bottom, then runs a valid conv."""
    k = w.shape[-1]
    total = k - 1
    lo, hi = total // 2, total - total // 2
    return F.conv2d(F.pad(x, (lo, hi, lo, hi)), w)


def generator_forward(state: State, noise: Tensor, training: bool) -> Tensor:
    """WGanGenerator.forward (generator.py:41-49): Upsample(2) -> [Upsample(2), conv 4x4 'same', BN, ReLU] x (L+1)
    -> Upsample(2) -> conv 4x4 'same' -> tanh.  ``noise_to_image`` indices: 0 = Upsample, 1..L+1 = DeConvBlocks,
    L+2 = Upsample, L+3 = Conv2d, L+4 = Tanh."""
    x = F.interpolate(noise, scale_factor=2, mode="nearest")
    i = 1
    while f"noise_to_image.{i}.de_conv_block.1.weight" in state:
        p = f"noise_to_image.{i}.de_conv_block."
        x = F.interpolate(x, scale_factor=2, mode="nearest")        # architecture.py:47-48
        x = same_conv_even(x, state[p + "1.weight"])               # :50-55 (bias=False)
        x = batch_norm(x, state, p + "2.", training)               # :61-62
        x = act(x, "relu")                                         # :65
        i += 1
    x = F.interpolate(x, scale_factor=2, mode="nearest")
    x = same_conv_even(x, state[f"noise_to_image.{i + 1}.weight"])
    return torch.tanh(x)


def discriminator_forward(state: State, x: Tensor, training: bool) -> Tensor:
    """WGanDiscriminator.forward (discriminator.py:40-46): 7x7 s2 reflect conv + BN + ReLU, MaxPool 3/2/1,
    [3x3 s2 zero-pad conv + BN + ReLU] x L, global average pool, Linear."""
    p = "feature_extractor.0.conv_block."
    x = act(batch_norm(conv2d(x, state[p + "0.weight"], None, 2, 3, "reflect"), state, p + "1.", training), "relu")
    x = F.max_pool2d(x, 3, 2, 1)
    i = 2
    while f"feature_extractor.{i}.conv_block.0.weight" in state:
        p = f"feature_extractor.{i}.conv_block."
        x = act(batch_norm(conv2d(x, state[p + "0.weight"], None, 2, 1, "zeros"), state, p + "1.", training), "relu")
        i += 1
    feat = x.mean(dim=(2, 3))
    return F.linear(feat, state["classifier.0.weight"], state["classifier.0.bias"])


class RMSprop:
    """torch.optim.RMSprop(lr) defaults: alpha 0.99, eps 1e-8, no momentum, not centered (base_trainer.py:72-73)."""

    def __init__(self, params: Sequence[Tensor], lr: float, alpha: float = 0.99, eps: float = 1e-8):
        self.params, self.lr, self.alpha, self.eps = list(params), lr, alpha, eps
        self.sq: List[Tensor] = [torch.zeros_like(p) for p in self.params]

    def zero_grad(self):
        for p in self.params:
            p.grad = None

    @torch.no_grad()
    def step(self):
        for p, sq in zip(self.params, self.sq):
            if p.grad is None:
                continue
            sq.mul_(self.alpha).addcmul_(p.grad, p.grad, value=1 - self.alpha)
            p.addcdiv_(p.grad, sq.sqrt().add_(self.eps), value=-self.lr)


class Trainer:
    """WGanTrainer._train_discriminator_once / _train_generator_once (wgan_trainer.py:124-138 / 115-122) with
    WGanModel.weight_clipping (wgan_model.py:18-21).  ``noise`` replaces the reference's ``torch.rand(B, nz, 1, 1)``
    draw (generator.py:46) so that both sides of a parity test see the same latent."""

    def __init__(self, g: State, d: State, lr: float = 5e-5, clipping_limit: float = 0.01):
        self.g, self.d, self.clip = g, d, clipping_limit
        for st in (g, d):
            for k in trainable_keys(st):
                st[k].requires_grad_(True)
        self.opt_g = RMSprop([g[k] for k in trainable_keys(g)], lr)
        self.opt_d = RMSprop([d[k] for k in trainable_keys(d)], lr)

    def train_discriminator_once(self, real: Tensor, noise: Tensor):
        with torch.no_grad():
            for k in trainable_keys(self.d):
                self.d[k].clamp_(-self.clip, self.clip)
        self.opt_d.zero_grad()
        fake = generator_forward(self.g, noise, True).detach()
        fake_logits = discriminator_forward(self.d, fake, True)
        real_logits = discriminator_forward(self.d, real, True)
        d_loss = -(real_logits.mean() - fake_logits.mean())
        d_loss.backward()
        self.opt_d.step()
        return d_loss.item()

    def train_generator_once(self, noise: Tensor):
        self.opt_g.zero_grad()
        for k in trainable_keys(self.d):  # the reference leaves D grads accumulating here; they are zeroed by the next D-step
            self.d[k].grad = None
        fake = generator_forward(self.g, noise, True)
        g_loss = -discriminator_forward(self.d, fake, True).mean()
        g_loss.backward()
        self.opt_g.step()
        return g_loss.item()
