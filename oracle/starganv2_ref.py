"""TEST INFRASTRUCTURE ONLY - plain-torch FP32 CPU restatement of the StarGAN-v2 path (BASELINE.json configs[3];
``norm_type='adain'``, ``w_hpf=0``) over reference-format ``state_dict``s of ``Generator`` / ``MappingNetwork`` /
``StyleEncoder`` / ``Discriminator`` (/root/reference/stargan-v2/core/model.py) and of the loss functions
(stargan-v2/core/solver.py:467-583, core/utils.py:485-490).  Pinned against the unmodified reference modules in
tests/test_oracle_starganv2.py (live, build container) and tests/golden/starganv2_small.pt.  Never imported by the product.
"""
from __future__ import annotations

import math
from types import SimpleNamespace

import torch
import torch.nn.functional as F

from .defectgan_ref import State, Tensor


def _conv(st: State, p: str, x: Tensor, pad: int) -> Tensor:
    return F.conv2d(x, st[p + "weight"], st.get(p + "bias"), padding=pad)


def _in(x: Tensor, eps=1e-5) -> Tensor:
    m = x.mean(dim=(2, 3), keepdim=True)
    v = x.var(dim=(2, 3), unbiased=False, keepdim=True)
    return (x - m) * torch.rsqrt(v + eps)


def _in_affine(st: State, p: str, x: Tensor) -> Tensor:
    return _in(x) * st[p + "weight"][None, :, None, None] + st[p + "bias"][None, :, None, None]


def res_blk(st: State, p: str, x: Tensor, downsample: bool) -> Tensor:
    """ResBlk.forward, model.py:45-68."""
    normalize = p + "norm1.weight" in st
    sc = x
    if p + "conv1x1.weight" in st:
        sc = _conv(st, p + "conv1x1.", sc, 0)
    if downsample:
        sc = F.avg_pool2d(sc, 2)
    h = _in_affine(st, p + "norm1.", x) if normalize else x
    h = _conv(st, p + "conv1.", F.leaky_relu(h, 0.2), 1)
    if downsample:
        h = F.avg_pool2d(h, 2)
    h = _in_affine(st, p + "norm2.", h) if normalize else h
    h = _conv(st, p + "conv2.", F.leaky_relu(h, 0.2), 1)
    return (sc + h) / math.sqrt(2)


def adain(st: State, p: str, x: Tensor, s: Tensor) -> Tensor:
    """AdaIN.forward, model.py:80-84."""
    h = F.linear(s, st[p + "fc.weight"], st[p + "fc.bias"])
    gamma, beta = torch.chunk(h[:, :, None, None], 2, dim=1)
    return (1 + gamma) * _in(x) + beta


def adain_res_blk(st: State, p: str, x: Tensor, s: Tensor, upsample: bool) -> Tensor:
    """AdainResBlk.forward (w_hpf = 0), model.py:100-123."""
    sc = F.interpolate(x, scale_factor=2, mode="nearest") if upsample else x
    if p + "conv1x1.weight" in st:
        sc = _conv(st, p + "conv1x1.", sc, 0)
    h = F.leaky_relu(adain(st, p + "norm1.", x, s), 0.2)
    if upsample:
        h = F.interpolate(h, scale_factor=2, mode="nearest")
    h = _conv(st, p + "conv1.", h, 1)
    h = F.leaky_relu(adain(st, p + "norm2.", h, s), 0.2)
    h = _conv(st, p + "conv2.", h, 1)
    return (h + sc) / math.sqrt(2)


def _count(st: State, stem: str) -> int:
    n = 0
    while any(k.startswith(f"{stem}.{n}.") for k in st):
        n += 1
    return n


def generator_forward(st: State, x: Tensor, s: Tensor) -> Tensor:
    """Generator.forward, model.py:371-393; the last two encode / first two decode blocks are the bottleneck (no resampling)."""
    n = _count(st, "encode")
    h = _conv(st, "from_rgb.", x, 1)
    for i in range(n):
        h = res_blk(st, f"encode.{i}.", h, downsample=i < n - 2)
    for i in range(n):
        h = adain_res_blk(st, f"decode.{i}.", h, s, upsample=i >= 2)
    h = F.leaky_relu(_in_affine(st, "to_rgb.0.", h), 0.2)
    return _conv(st, "to_rgb.2.", h, 0)


def _mlp(st: State, p: str, h: Tensor) -> Tensor:
    i = 0
    while f"{p}{i}.weight" in st:
        h = F.linear(h, st[f"{p}{i}.weight"], st[f"{p}{i}.bias"])
        if f"{p}{i + 2}.weight" in st:
            h = F.relu(h)
        i += 2
    return h


def mapping_forward(st: State, z: Tensor, y: Tensor) -> Tensor:
    """MappingNetwork.forward, model.py:462-471."""
    h = F.relu(_mlp(st, "shared.", z))  # the shared stack ends with a ReLU (model.py:448-451)
    out = torch.stack([_mlp(st, f"unshared.{d}.", h) for d in range(_count(st, "unshared"))], dim=1)
    return out[torch.arange(y.size(0)), y]


def _trunk(st: State, p: str, x: Tensor) -> Tensor:
    n = _count(st, p.rstrip("."))
    blocks = [i for i in range(1, n) if f"{p}{i}.conv1.weight" in st]
    h = _conv(st, f"{p}0.", x, 1)
    for i in blocks:
        h = res_blk(st, f"{p}{i}.", h, downsample=True)
    last = blocks[-1]
    h = F.leaky_relu(h, 0.2)
    h = _conv(st, f"{p}{last + 2}.", h, 0)
    return F.leaky_relu(h, 0.2), last


def style_encoder_forward(st: State, x: Tensor, y: Tensor) -> Tensor:
    """StyleEncoder.forward, model.py:494-503."""
    h = _trunk(st, "shared.", x)[0].flatten(1)
    out = torch.stack([F.linear(h, st[f"unshared.{d}.weight"], st[f"unshared.{d}.bias"]) for d in range(_count(st, "unshared"))], dim=1)
    return out[torch.arange(y.size(0)), y]


def discriminator_forward(st: State, x: Tensor, y: Tensor) -> Tensor:
    """Discriminator.forward, model.py:526-532."""
    h, last = _trunk(st, "main.", x)
    out = _conv(st, f"main.{last + 4}.", h, 0).flatten(1)
    return out[torch.arange(y.size(0)), y]


def adv_loss(logits: Tensor, target: int) -> Tensor:
    return F.binary_cross_entropy_with_logits(logits, torch.full_like(logits, fill_value=target))


def r1_reg(d_out: Tensor, x_in: Tensor) -> Tensor:
    """solver.py:573-583."""
    (g,) = torch.autograd.grad(d_out.sum(), x_in, create_graph=True, retain_graph=True)
    return 0.5 * g.pow(2).view(x_in.size(0), -1).sum(1).mean(0)


def style_code(nets, y, x_ref, z):
    return mapping_forward(nets.mapping_network, z, y) if z is not None else style_encoder_forward(nets.style_encoder, x_ref, y)


def compute_d_loss(nets, args, x_real, y_org, y_trg, z_trg=None, x_ref=None):
    """solver.py:467-491; ``nets``: namespace of state dicts."""
    x_real = x_real.clone().requires_grad_(True)
    out = discriminator_forward(nets.discriminator, x_real, y_org)
    loss_real = adv_loss(out, 1)
    loss_reg = r1_reg(out, x_real)
    with torch.no_grad():
        x_fake = generator_forward(nets.generator, x_real, style_code(nets, y_trg, x_ref, z_trg))
    loss_fake = adv_loss(discriminator_forward(nets.discriminator, x_fake, y_trg), 0)
    loss = loss_real + loss_fake + args.lambda_reg * loss_reg
    return loss, SimpleNamespace(real=loss_real.item(), fake=loss_fake.item(), reg=loss_reg.item())


def compute_g_loss(nets, args, x_real, y_org, y_trg, z_trgs=None, x_refs=None):
    """solver.py:494-546."""
    z_trg, z_trg2 = z_trgs if z_trgs is not None else (None, None)
    x_ref, x_ref2 = x_refs if x_refs is not None else (None, None)
    s_trg = style_code(nets, y_trg, x_ref, z_trg)
    x_fake = generator_forward(nets.generator, x_real, s_trg)
    loss_adv = adv_loss(discriminator_forward(nets.discriminator, x_fake, y_trg), 1)
    s_pred = style_encoder_forward(nets.style_encoder, x_fake, y_trg)
    loss_sty = torch.mean(torch.abs(s_pred - s_trg))
    s_trg2 = style_code(nets, y_trg, x_ref2, z_trg2)
    x_fake2 = generator_forward(nets.generator, x_real, s_trg2).detach()
    loss_ds = torch.mean(torch.abs(x_fake - x_fake2))
    s_org = style_encoder_forward(nets.style_encoder, x_real, y_org)
    x_rec = generator_forward(nets.generator, x_fake, s_org)
    loss_cyc = torch.mean(torch.abs(x_rec - x_real))
    loss = loss_adv + args.lambda_sty * loss_sty - args.lambda_ds * loss_ds + args.lambda_cyc * loss_cyc
    return loss, SimpleNamespace(adv=loss_adv.item(), sty=loss_sty.item(), ds=loss_ds.item(), cyc=loss_cyc.item())


def mapping_state(seed: int, latent_dim: int, style_dim: int, num_domains: int) -> State:
    """Deterministic He-initialised MappingNetwork state dict (its hard-coded 512-wide layers are 3 M parameters: the golden
    fixture stores the seed and a checksum instead of the tensors)."""
    g = torch.Generator().manual_seed(seed)
    st: State = {}

    def lin(p, i, o):
        st[p + "weight"] = torch.randn(o, i, generator=g) * math.sqrt(2.0 / i)
        st[p + "bias"] = torch.zeros(o)

    lin("shared.0.", latent_dim, 512)
    for j in (2, 4, 6):
        lin(f"shared.{j}.", 512, 512)
    for d in range(num_domains):
        for j in (0, 2, 4):
            lin(f"unshared.{d}.{j}.", 512, 512)
        lin(f"unshared.{d}.6.", 512, style_dim)
    return st
