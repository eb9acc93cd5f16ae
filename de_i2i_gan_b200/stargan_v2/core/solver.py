"""StarGAN-v2 loss functions (stargan-v2/core/solver.py:467-583, core/utils.py:485-495) on the CUDA path: same names,
arguments and returned (loss, Munch-like log) pairs.  ``Solver.train`` itself is broken in the reference as checked in
(SURVEY.md 0.8); ``train_iteration`` below is the four sub-steps of its loop body (solver.py:258-308) with FlatAdam."""
from types import SimpleNamespace

import torch

from ... import functional as Fn
from ... import ops


class _DeferredLog(SimpleNamespace):
    """Loss scalars still on the device.  The reference's loss functions return python floats (`.item()`, solver.py:491,546): fourteen
    host synchronisations per iteration, each of which drains the GPU before the next kernel can be queued (and none of which a
    CUDA graph can contain).  Called on their own, compute_d_loss / compute_g_loss still do exactly that; `train_iteration` asks for
    the deferred form and reads all fourteen scalars with ONE transfer after the last sub-step has been queued."""


_defer_logs = False


def _log(**scalars):
    if _defer_logs:
        return _DeferredLog(**{k: v.detach().float().reshape(()) for k, v in scalars.items()})
    return SimpleNamespace(**{k: v.item() for k, v in scalars.items()})


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
    return loss, _log(real=loss_real, fake=loss_fake, reg=loss_reg)


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
    lambda_ds = getattr(args, "_lambda_ds_dev", None)  # device copy while train_iteration runs (its decay must reach captured graphs)
    if lambda_ds is None:
        lambda_ds = args.lambda_ds
    loss = loss_adv + args.lambda_sty * loss_sty - lambda_ds * loss_ds + args.lambda_cyc * loss_cyc
    return loss, _log(adv=loss_adv, sty=loss_sty, ds=loss_ds, cyc=loss_cyc)


@torch.no_grad()
def moving_average(model, model_test, beta=0.999):
    """solver.py:549-551: param_test = lerp(param, param_test, beta) (dei2i_ema, in place)."""
    for param, param_test in zip(model.parameters(), model_test.parameters()):
        ops.ema_(param.data.contiguous(), param_test.data, beta)


# ------------------------------------------------------------------------------------------------ training iteration
def he_init(module):
    """core/utils.py:38-46: Kaiming-normal weights, zero bias, for every nn.Conv2d / nn.Linear."""
    if isinstance(module, (torch.nn.Conv2d, torch.nn.Linear)):
        torch.nn.init.kaiming_normal_(module.weight, mode="fan_in", nonlinearity="relu")
        if module.bias is not None:
            torch.nn.init.constant_(module.bias, 0)


def build_training_state(args, device):
    """What ``Solver.__init__`` sets up for training in adain mode (solver.py:49-91): the four networks + their EMA copies
    (``build_model``), he_init of every network, and one Adam per network (lr ``args.lr``, the mapping network ``args.f_lr``,
    betas (beta1, beta2), weight decay) - FlatAdam on the fused kernel.  Returns a namespace (nets, nets_ema, optims,
    initial_lambda_ds)."""
    from ...optim import FlatAdam
    from . import model as M
    nets, nets_ema = M.build_nets(args)
    for name, net in vars(nets).items():
        net.apply(he_init)
        net.to(device)
    for name, net in vars(nets_ema).items():
        net.load_state_dict(getattr(nets, name).state_dict())  # solver.py: the EMA copies start as deep copies of the live nets
        net.to(device)
    optims = SimpleNamespace(**{
        name: FlatAdam(list(net.parameters()), lr=getattr(args, "f_lr", 1e-6) if name == "mapping_network" else getattr(args, "lr", 1e-4),
                       betas=(getattr(args, "beta1", 0.0), getattr(args, "beta2", 0.99)), weight_decay=getattr(args, "weight_decay", 1e-4))
        for name, net in vars(nets).items()})
    return SimpleNamespace(nets=nets, nets_ema=nets_ema, optims=optims, initial_lambda_ds=args.lambda_ds)


def _reset_grad(optims):
    """solver.py:127-129."""
    for o in vars(optims).values():
        o.zero_grad()


def train_iteration(state, args, inputs, masks=None):
    """One pass of the loop body of ``Solver.train`` in adain mode (solver.py:258-308): discriminator on latent-generated and
    on reference-generated fakes (each with the R1 double backward), generator on latent codes (+ mapping network and style
    encoder steps) and on reference images, the three EMA updates, and the linear decay of the diversity-loss weight.
    ``inputs``: namespace with x_src, y_src, x_ref, x_ref2, y_ref, z_trg, z_trg2 (what InputFetcher yields).  Returns the four log
    namespaces (d_latent, d_ref, g_latent, g_ref) of python floats.

    Each of the four sub-steps (loss graph incl. the double backward, zero_grad, backward, its Adam steps, the EMA updates) is one
    CUDA graph per input-shape signature (GraphedStep: call 1 eager, call 2 captures, then replays; DEI2I_CUDA_GRAPH=0 disables):
    an iteration is ~5700 kernel launches, which eagerly is bound by the host.  The diversity weight is read from device memory so
    that its decay reaches the captured graphs; the fourteen loss scalars come back in one transfer at the end."""
    global _defer_logs
    x_real, y_org = inputs.x_src, inputs.y_src
    x_ref, x_ref2, y_trg = inputs.x_ref, inputs.x_ref2, inputs.y_ref
    z_trg, z_trg2 = inputs.z_trg, inputs.z_trg2
    steps = _sub_steps(state, args, masks)
    state.lambda_ds_dev.fill_(float(args.lambda_ds))  # kernel argument, no synchronisation
    args._lambda_ds_dev = state.lambda_ds_dev
    _defer_logs = True
    try:
        outs = [steps["d_latent"](x_real, y_org, y_trg, z_trg),
                steps["d_ref"](x_real, y_org, y_trg, x_ref),
                steps["g_latent"](x_real, y_org, y_trg, z_trg, z_trg2),
                steps["g_ref"](x_real, y_org, y_trg, x_ref, x_ref2)]
    finally:
        _defer_logs = False
        args._lambda_ds_dev = None
    if args.lambda_ds > 0:  # solver.py:303-304
        args.lambda_ds -= state.initial_lambda_ds / args.ds_iter
    flat = torch.cat(outs).tolist()  # the iteration's one device-to-host read
    d_keys, g_keys = ("real", "fake", "reg"), ("adv", "sty", "ds", "cyc")
    logs, i = [], 0
    for keys in (d_keys, d_keys, g_keys, g_keys):
        logs.append(SimpleNamespace(**dict(zip(keys, flat[i:i + len(keys)]))))
        i += len(keys)
    return tuple(logs)


def _sub_steps(state, args, masks):
    """The four sub-step bodies of an iteration behind GraphedStep wrappers, built once per training state."""
    cached = getattr(state, "_sub_steps", None)
    if cached is not None and cached[0] is args and cached[1] is masks:
        return cached[2]
    import os
    from ...trainers.defectgan_trainer import GraphedStep
    nets, nets_ema, optims = state.nets, state.nets_ema, state.optims
    dev = next(nets.generator.parameters()).device
    if getattr(state, "lambda_ds_dev", None) is None:
        state.lambda_ds_dev = torch.zeros((), dtype=torch.float32, device=dev)

    def d_body(kind):
        def body(x_real, y_org, y_trg, aux):
            with Fn.weight_cache():
                # compute_d_loss marks its image argument as requiring grad (R1); a detached alias keeps that off ``x_real`` itself
                d_loss, log = compute_d_loss(nets, args, x_real.detach(), y_org, y_trg, masks=masks, **{kind: aux})
                _reset_grad(optims)
                d_loss.backward()
            optims.discriminator.step()
            return torch.stack(list(vars(log).values()))
        return body

    def g_latent(x_real, y_org, y_trg, z_trg, z_trg2):
        with Fn.weight_cache():
            g_loss, log = compute_g_loss(nets, args, x_real, y_org, y_trg, z_trgs=[z_trg, z_trg2], masks=masks)
            _reset_grad(optims)
            g_loss.backward()
        optims.generator.step()
        optims.mapping_network.step()
        optims.style_encoder.step()
        return torch.stack(list(vars(log).values()))

    def g_ref(x_real, y_org, y_trg, x_ref, x_ref2):
        with Fn.weight_cache():
            g_loss, log = compute_g_loss(nets, args, x_real, y_org, y_trg, x_refs=[x_ref, x_ref2], masks=masks)
            _reset_grad(optims)
            g_loss.backward()
        optims.generator.step()
        moving_average(nets.generator, nets_ema.generator, beta=0.999)
        moving_average(nets.mapping_network, nets_ema.mapping_network, beta=0.999)
        moving_average(nets.style_encoder, nets_ema.style_encoder, beta=0.999)
        return torch.stack(list(vars(log).values()))

    use_graph = os.environ.get("DEI2I_CUDA_GRAPH", "1") != "0" and dev.type == "cuda" and masks is None
    pool, staged = [None], {}
    steps = {"d_latent": GraphedStep(d_body("z_trg"), optims.discriminator, use_graph, pool, staged),
             "d_ref": GraphedStep(d_body("x_ref"), optims.discriminator, use_graph, pool, staged),
             "g_latent": GraphedStep(g_latent, [optims.generator, optims.mapping_network, optims.style_encoder], use_graph, pool, staged),
             "g_ref": GraphedStep(g_ref, optims.generator, use_graph, pool, staged)}
    state._sub_steps = (args, masks, steps)
    return steps
