"""WGAN path (BASELINE.json configs[2]) on the CUDA kernels against tests/golden/wgan_small.pt (unmodified reference
networks on CPU, FP32; oracle/make_golden_wgan.py).  BF16 tolerances as in tests/test_gpu_golden.py."""
import os
from types import SimpleNamespace

import pytest
import torch
import torch.nn.functional as F

from conftest import rel_err, GOLDEN

pytestmark = pytest.mark.gpu
TOL_ACT = 2e-2


def _load():
    return torch.load(os.path.join(GOLDEN, "wgan_small.pt"), weights_only=False)


def _nets(t):
    from de_i2i_gan_b200.models.networks.discriminator import WGanDiscriminator
    from de_i2i_gan_b200.models.networks.generator import WGanGenerator
    c = t["cfg"]
    L = c["image_size"].bit_length() - 1 - 3
    G = WGanGenerator(L, ngf=c["ngf"], noise_dim=c["noise_dim"]).cuda()
    D = WGanDiscriminator(L, ndf=c["ndf"]).cuda()
    assert list(G.state_dict()) == list(t["g_state"]) and list(D.state_dict()) == list(t["d_state"])  # drop-in checkpoints
    G.load_state_dict(t["g_state"])
    D.load_state_dict(t["d_state"])
    return G, D


def _cos(a, b):
    a, b = a.detach().float().cpu().flatten(), b.detach().float().cpu().flatten()
    return (torch.dot(a, b) / (a.norm() * b.norm() + 1e-30)).item()


def test_wgan_generator_and_critic_forward_backward():
    t = _load()
    G, D = _nets(t)
    G.train(); D.train()
    fake = G(t["z"].cuda())
    assert rel_err(fake, t["fwd"]["fake"]) < TOL_ACT
    fake.square().mean().backward()
    got = torch.cat([p.grad.flatten().cpu() for _, p in G.named_parameters()])
    ref = torch.cat([t["fwd"]["g_grads"][n].flatten() for n, _ in G.named_parameters()])
    assert ((got - ref).norm() / ref.norm()).item() < 8e-2
    x = t["real"].cuda().requires_grad_(True)
    logits = D(x)
    assert tuple(logits.shape) == tuple(t["fwd"]["logits"].shape)
    assert rel_err(logits, t["fwd"]["logits"]) < TOL_ACT
    logits.sum().backward()
    assert _cos(x.grad, t["fwd"]["dx"]) > 0.98
    got = torch.cat([p.grad.flatten().cpu() for _, p in D.named_parameters()])
    ref = torch.cat([t["fwd"]["d_grads"][n].flatten() for n, _ in D.named_parameters()])
    assert ((got - ref).norm() / ref.norm()).item() < 8e-2
    G.eval(); D.eval()
    with torch.no_grad():
        assert rel_err(G(t["z"].cuda()), t["fwd"]["fake_eval"]) < TOL_ACT
        # eval-mode logits are ~1e-4: a cancelling sum of 32 pooled features x N(0, 0.02) weights -> twice the BF16 budget
        assert rel_err(D(t["real"].cuda()), t["fwd"]["logits_eval"]) < 2 * TOL_ACT


def test_wgan_two_training_steps():
    from de_i2i_gan_b200 import options
    from de_i2i_gan_b200.trainers.wgan_trainer import WGanTrainer
    t = _load()
    c = t["cfg"]
    import pathlib
    opt = options.wgan_options(image_size=c["image_size"], ngf=c["ngf"], ndf=c["ndf"], noise_dim=c["noise_dim"], batch_size=c["batch"],
                               lr=c["lr"], clipping_limit=c["clipping_limit"], num_display_images=4,
                               ckpt_dir=pathlib.Path("/tmp/dei2i_wgan_ckpt"))
    tr = WGanTrainer(opt)
    tr.model.netG.load_state_dict(t["g_state"])
    tr.model.netD.load_state_dict(t["d_state"])
    real = t["real"].cuda()
    for it in range(2):
        tr._train_discriminator_once(real, noise=t["zs"][2 * it].cuda())
        tr._train_generator_once(t["zs"][2 * it + 1].cuda())
        ref = t["steps"][it]
        # the losses are differences of O(1e-4) critic outputs: compare on the scale of the critic's logits
        scale = max(t["fwd"]["logits"].abs().max().item(), 1e-3)
        assert abs(tr.losses["gan_D"][-1] - ref["d_loss"]) < TOL_ACT * scale, (it, tr.losses["gan_D"][-1], ref)
        assert abs(tr.losses["gan_G"][-1] - ref["g_loss"]) < TOL_ACT * scale, (it, tr.losses["gan_G"][-1], ref)
    for net, before, after in ((tr.model.netG, t["g_state"], t["g_state_after"]), (tr.model.netD, t["d_state"], t["d_state_after"])):
        st = net.state_dict()
        keys = [k for k, _ in net.named_parameters()]
        d_ref = torch.cat([(after[k] - before[k]).flatten() for k in keys])
        d_got = torch.cat([(st[k].cpu() - before[k]).flatten() for k in keys])
        # RMSprop's first steps move EVERY element by ~10 * lr * sign(g) (v = 0.01 g^2), so an element whose gradient is at
        # the BF16 noise level counts as much as a dominant one: sign agreement, not magnitude, is what this cosine measures.
        # The kernel itself is checked exactly in test_rmsprop_clamp_ema_kernels, the gradients in the test above.
        assert _cos(d_got, d_ref) > 0.5
        for k in st:
            if k.endswith("running_var"):
                assert rel_err(st[k], after[k]) < 5e-2, k
            # running means are not compared: RMSprop's sign-like first steps (see above) change each weight by 2.5 %, and the
            # ~0 batch means of zero-mean layers move with them


@pytest.mark.parametrize("k,s,p,h,w", [(3, 2, 1, 16, 16), (3, 2, 1, 9, 7), (2, 2, 0, 8, 8)])
def test_maxpool_fwd_bwd(k, s, p, h, w):
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(1)
    x = F.relu(torch.randn(2, 16, h, w, generator=g)).to(torch.bfloat16).float().requires_grad_(True)  # ReLU zeros -> ties
    y = F.max_pool2d(x, k, s, p)
    dy = torch.randn(y.shape, generator=g).to(torch.bfloat16).float()
    (dx_ref,) = torch.autograd.grad(y, x, dy)
    xd = x.detach().permute(0, 2, 3, 1).contiguous().to(torch.bfloat16).cuda()
    out, idx = ops.maxpool_fwd(xd, k, s, p)
    assert torch.equal(out.float().cpu().permute(0, 3, 1, 2), y.detach())
    dx = ops.maxpool_bwd(dy.permute(0, 2, 3, 1).contiguous().to(torch.bfloat16).cuda(), idx, (h, w), k, s, p)
    assert rel_err(dx.float().cpu().permute(0, 3, 1, 2), dx_ref) < 1e-2


def test_upsample_pad_avgpool_gap():
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(2)
    x = torch.randn(2, 8, 5, 6, generator=g).to(torch.bfloat16).float().requires_grad_(True)
    ref = F.pad(F.interpolate(x, scale_factor=2, mode="nearest"), (1, 2, 1, 2))
    dy = torch.randn(ref.shape, generator=g).to(torch.bfloat16).float()
    (dx_ref,) = torch.autograd.grad(ref, x, dy)
    to = lambda t_: t_.detach().permute(0, 2, 3, 1).contiguous().to(torch.bfloat16).cuda()
    back = lambda t_: t_.float().cpu().permute(0, 3, 1, 2)
    out = ops.upsample_pad(to(x), 2, (1, 1, 2, 2))
    assert torch.equal(back(out), ref.detach())
    assert rel_err(back(ops.upsample_pad_bwd(to(dy), (5, 6), 2, (1, 1, 2, 2))), dx_ref) < 1e-2
    x2 = torch.randn(2, 16, 8, 12, generator=g).to(torch.bfloat16).float()
    assert rel_err(back(ops.avgpool_fwd(to(x2), 2, 0.5)), F.avg_pool2d(x2, 2) * 0.5) < 1e-2
    d2 = torch.randn(2, 16, 4, 6, generator=g).to(torch.bfloat16).float()
    assert rel_err(back(ops.avgpool_bwd(to(d2), 2, 0.5)), F.interpolate(d2, scale_factor=2) * 0.125) < 1e-2
    assert rel_err(ops.gap_fwd(to(x2)).cpu(), x2.mean(dim=(2, 3))) < 1e-3
    dg = torch.randn(2, 16, generator=g)
    assert rel_err(back(ops.gap_bwd(dg.cuda(), (8, 12))), (dg / 96)[:, :, None, None].expand(2, 16, 8, 12)) < 1e-2


def test_rmsprop_clamp_ema_kernels():
    """dei2i_rmsprop / dei2i_clamp / dei2i_ema against torch.optim.RMSprop, Tensor.clamp_, torch.lerp (FP32, exact to rounding)."""
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(3)
    p0 = torch.randn(10007, generator=g)
    ref = p0.clone().requires_grad_(True)
    opt = torch.optim.RMSprop([ref], lr=5e-5)
    p, sq = p0.clone().cuda(), torch.zeros(10007, device="cuda")
    for _ in range(3):
        grad = torch.randn(10007, generator=g) * 1e-3
        ref.grad = grad.clone()
        opt.step()
        ops.rmsprop(p, grad.cuda(), sq, 5e-5)
    assert rel_err(p, ref) < 1e-6
    ops.clamp_(p, -0.03, 0.03)
    assert torch.equal(p.cpu(), p.cpu().clamp(-0.03, 0.03)) and rel_err(p, ref.detach().clamp(-0.03, 0.03)) < 1e-6
    a, b = torch.randn(4099, generator=g), torch.randn(4099, generator=g)
    bd = b.clone().cuda()
    ops.ema_(a.cuda(), bd, 0.999)
    assert rel_err(bd, torch.lerp(a, b, 0.999)) < 1e-6  # moving_average, stargan-v2/core/solver.py:549-551
