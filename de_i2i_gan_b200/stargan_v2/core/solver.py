"""StarGAN-v2 loss functions (stargan-v2/core/solver.py:467-583, core/utils.py:485-495) on the CUDA path: same names,
arguments and returned (loss, Munch-like log) pairs.  ``Solver.train`` itself is broken in the reference as checked in
(SURVEY.md 0.8); ``train_iteration`` below is the four sub-steps of its loop body (solver.py:258-308) with FlatAdam."""
from types import SimpleNamespace

import torch

from ... import functional as Fn
from ... import ops


def get_style_code(nets, norm_type, num_embeds, y_trg, x_ref, z_trg):
    """core/utils.py:485-490 (adain branch)."""
    assert norm_type == "adain"
    if z_trg is not None:
        return nets.mapping_network(z_trg, y_trg)
    return nets.style_encoder(x_ref, y_trg)


def adv_loss(logits, target):
    """solver.py:566-570: BCE-with-logits against an all-``target`` tensor (fused loss + gradient kernel)."""
    assert target in [1, 0]
    return Fn.bce_with_logits(logits, float(target))


def r1_reg(d_out, x_in):
    """solver.py:573-583: zero-centred gradient penalty; the double backward runs through Fn._DgradL / _ActBwd / _AvgPoolBwd."""
    batch_size = x_in.size(0)
    with Fn.input_grad_only():  # this backward only needs d(out)/d(x): skip the weight / bias gradient kernels
        grad_dout = torch.autograd.grad(outputs=d_out.sum(), inputs=x_in, create_graph=True, retain_graph=True, only_inputs=True)[0]
    grad_dout2 = grad_dout.pow(2)
    assert grad_dout2.size() == x_in.size()
    return 0.5 * grad_dout2.view(batch_size, -1).sum(1).mean(0)


def compute_d_loss(nets, args, x_real, y_org, y_trg, z_trg=None, x_ref=None, masks=None):
    """solver.py:467-491."""
    assert (z_trg is None) != (x_ref is None)
    x_real.requires_grad_()
    out = nets.discriminator(x_real, y_org)
    loss_real = adv_loss(out, 1)
    loss_reg = r1_reg(out, x_real)
    with torch.no_grad():
        s_trg = get_style_code(nets, args.norm_type, 1, y_trg, x_ref, z_trg)
        x_fake = nets.generator(x_real, s_trg, labels=y_trg, masks=masks)
    out = nets.discriminator(x_fake, y_trg)
    loss_fake = adv_loss(out, 0)
    loss = loss_real + loss_fake + args.lambda_reg * loss_reg
    return loss, SimpleNamespace(real=loss_real.item(), fake=loss_fake.item(), reg=loss_reg.item())


def compute_g_loss(nets, args, x_real, y_org, y_trg, z_trgs=None, x_refs=None, masks=None):
    """solver.py:494-546."""
    assert (z_trgs is None) != (x_refs is None)
    z_trg, z_trg2 = z_trgs if z_trgs is not None else (None, None)
    x_ref, x_ref2 = x_refs if x_refs is not None else (None, None)
    s_trg = get_style_code(nets, args.norm_type, args.num_embeds, y_trg, x_ref, z_trg)
    x_fake = nets.generator(x_real, s_trg, labels=y_trg, masks=masks)
    out = nets.discriminator(x_fake, y_trg)
    loss_adv = adv_loss(out, 1)
    s_pred = get_style_code(nets, args.norm_type, args.num_embeds, y_trg, x_fake, z_trg=None)
    loss_sty = Fn.l1_loss(s_pred, s_trg)
    s_trg2 = get_style_code(nets, args.norm_type, args.num_embeds, y_trg, x_ref2, z_trg=z_trg2)
    x_fake2 = nets.generator(x_real, s_trg2, labels=y_trg, masks=masks).detach()
    loss_ds = Fn.l1_loss(x_fake, x_fake2)
    s_org = get_style_code(nets, args.norm_type, args.num_embeds, y_org, x_real, z_trg=None)
    x_rec = nets.generator(x_fake, s_org, labels=y_org, masks=None)
    loss_cyc = Fn.l1_loss(x_rec, x_real)
    loss = loss_adv + args.lambda_sty * loss_sty - args.lambda_ds * loss_ds + args.lambda_cyc * loss_cyc
    return loss, SimpleNamespace(adv=loss_adv.item(), sty=loss_sty.item(), ds=loss_ds.item(), cyc=loss_cyc.item())


@torch.no_grad()
def moving_average(model, model_test, beta=0.999):
    """solver.py:549-551: param_test = lerp(param, param_test, beta) (dei2i_ema, in place)."""
    for param, param_test in zip(model.parameters(), model_test.parameters()):
        ops.ema_(param.data.contiguous(), param_test.data, beta)
