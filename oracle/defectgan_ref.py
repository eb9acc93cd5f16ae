"""TEST INFRASTRUCTURE ONLY - plain-torch FP32 CPU restatement of the DefectGAN hot path.

Every function cites the reference file:line (relative to /root/reference/defectGAN) it follows.
The networks are re-stated *functionally* over a reference-format ``state_dict`` (same keys, shapes,
dtypes as ``DefectGanGenerator(opt).state_dict()`` / ``DefectGanDiscriminator(opt).state_dict()``),
so a reference checkpoint can be fed to the oracle and to the CUDA modules alike.

This restatement is pinned against the unmodified reference in tests/test_oracle_vs_reference.py (live,
build container only) and against tests/golden/*.pt (committed fixtures produced by oracle/make_golden.py
from the unmodified reference).  It is never imported by the product package.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor
State = Dict[str, Tensor]


# --------------------------------------------------------------------------------------------------
# leaf ops
# --------------------------------------------------------------------------------------------------
def _normalize(v: Tensor, eps: float) -> Tensor:
    # torch.nn.functional.normalize(v, dim=0, eps): v / max(||v||, eps)
    return v / v.norm().clamp_min(eps)


def spectral_weight(state: State, prefix: str, training: bool, eps: float = 1e-12) -> Tensor:
    """Old-hook-API ``torch.nn.utils.spectral_norm`` as applied at models/networks/architecture.py:68-71,
    109-112,238-239,338-341: one power iteration per *forward call* in training mode (u, v updated in
    place under no_grad, then cloned), sigma = u^T W v differentiable w.r.t. W only, weight = W / sigma."""
    w = state[prefix + "weight_orig"]
    u = state[prefix + "weight_u"]
    v = state[prefix + "weight_v"]
    wm = w.reshape(w.shape[0], -1)
    if training:
        with torch.no_grad():
            v.copy_(_normalize(torch.mv(wm.t(), u), eps))
            u.copy_(_normalize(torch.mv(wm, v), eps))
        u = u.clone()
        v = v.clone()
    sigma = torch.dot(u, torch.mv(wm, v))
    return w / sigma


def conv_weight(state: State, prefix: str, training: bool) -> Tensor:
    """Returns the effective conv weight for module ``prefix`` (spectral or plain)."""
    if prefix + "weight_orig" in state:
        return spectral_weight(state, prefix, training)
    return state[prefix + "weight"]


def conv2d(x: Tensor, w: Tensor, bias: Optional[Tensor] = None, stride: int = 1, pad: int = 0,
           pad_mode: str = "zeros") -> Tensor:
    """nn.Conv2d forward: with padding_mode != 'zeros' torch pads with F.pad(mode) then runs a valid conv
    (torch.nn.Conv2d._conv_forward); call sites architecture.py:51-56,95-100,228-233,320-337."""
    if pad > 0 and pad_mode == "reflect":
        x = F.pad(x, (pad, pad, pad, pad), mode="reflect")
        pad = 0
    return F.conv2d(x, w, bias, stride=stride, padding=pad)


def same_pad(k: int) -> int:
    return (k - 1) // 2


def batch_norm(x: Tensor, state: State, prefix: str, training: bool, momentum: float = 0.1,
               eps: float = 1e-5) -> Tensor:
    """nn.BatchNorm2d (generator.py:71,113,124 through ConvBlock architecture.py:103-104): batch statistics
    and running-stat update (unbiased var) in training mode, running statistics in eval mode."""
    w, b = state[prefix + "weight"], state[prefix + "bias"]
    rm, rv = state[prefix + "running_mean"], state[prefix + "running_var"]
    if training:
        mean = x.mean(dim=(0, 2, 3))
        var = x.var(dim=(0, 2, 3), unbiased=False)
        with torch.no_grad():
            n = x.numel() / x.shape[1]
            rm.mul_(1 - momentum).add_(momentum * mean.detach())
            rv.mul_(1 - momentum).add_(momentum * var.detach() * n / max(n - 1, 1))
            if prefix + "num_batches_tracked" in state:
                state[prefix + "num_batches_tracked"] += 1
    else:
        mean, var = rm, rv
    xh = (x - mean[None, :, None, None]) * torch.rsqrt(var[None, :, None, None] + eps)
    return xh * w[None, :, None, None] + b[None, :, None, None]


def instance_norm(x: Tensor, eps: float = 1e-5) -> Tensor:
    """nn.InstanceNorm2d(affine=False, track_running_stats=False) (normalization.py:14): per-(n,c)
    mean and biased variance over H*W."""
    mean = x.mean(dim=(2, 3), keepdim=True)
    var = x.var(dim=(2, 3), unbiased=False, keepdim=True)
    return (x - mean) * torch.rsqrt(var + eps)


def act(x: Tensor, kind: Optional[str]) -> Tensor:
    """get_act_layer, architecture.py:12-26."""
    if kind is None:
        return x
    if kind == "leaky_relu":
        return F.leaky_relu(x, 0.2)
    if kind == "relu":
        return F.relu(x)
    if kind == "tanh":
        return torch.tanh(x)
    if kind == "sigmoid":
        return torch.sigmoid(x)
    raise NameError(kind)


class NoiseSource:
    """Feeds NoiseInjection (architecture.py:385-389).  With ``noises=None`` draws
    ``image.new_empty(B,1,H,W).normal_()`` from the global torch RNG in call order exactly like the
    reference; otherwise pops pre-drawn tensors in call order.  ``drawn`` records what was used."""

    def __init__(self, noises: Optional[Sequence[Tensor]] = None):
        self.queue = list(noises) if noises is not None else None
        self.drawn: List[Tensor] = []

    def __call__(self, like: Tensor) -> Tensor:
        b, _, h, w = like.shape
        if self.queue is None:
            n = like.new_empty(b, 1, h, w).normal_()
        else:
            n = self.queue.pop(0)
            assert tuple(n.shape) == (b, 1, h, w), (n.shape, like.shape)
        self.drawn.append(n)
        return n


def noise_injection(x: Tensor, state: State, prefix: str, noise: NoiseSource) -> Tensor:
    """NoiseInjection.forward, architecture.py:385-389: image + weight * noise, weight of shape (1,1,1,1)."""
    if prefix + "weight" not in state:
        return x
    return x + state[prefix + "weight"] * noise(x)


# --------------------------------------------------------------------------------------------------
# blocks
# --------------------------------------------------------------------------------------------------
def conv_block(x: Tensor, state: State, prefix: str, k: int, stride: int, pad: int, pad_mode: str,
               norm: bool, act_kind: Optional[str], training: bool) -> Tensor:
    """ConvBlock, architecture.py:79-118: [spectral]Conv2d(bias=False) -> [BatchNorm2d] -> act.
    ``prefix`` ends with 'conv_block.'."""
    w = conv_weight(state, prefix + "0.", training)
    y = conv2d(x, w, state.get(prefix + "0.bias"), stride, pad, pad_mode)
    if norm:
        y = batch_norm(y, state, prefix + "1.", training)
    return act(y, act_kind)


def res_block(x: Tensor, state: State, prefix: str, training: bool) -> Tensor:
    """ResBlock (down_scale=False), architecture.py:121-176, as built at generator.py:118-126:
    x + CB(3x3 reflect, BN, -)(CB(3x3 reflect, BN, LReLU)(x))."""
    y = conv_block(x, state, prefix + "res_block.0.conv_block.", 3, 1, 1, "reflect", True, "leaky_relu", training)
    y = conv_block(y, state, prefix + "res_block.1.conv_block.", 3, 1, 1, "reflect", True, None, training)
    return y + x


def spade(x: Tensor, seg: Tensor, state: State, prefix: str) -> Tensor:
    """SPADE.forward, normalization.py:24-37 with norm_layer=InstanceNorm2d, 3x3 zero 'same' convs with bias."""
    normalized = instance_norm(x)
    seg = F.interpolate(seg, size=x.shape[2:], mode="nearest")
    actv = F.relu(F.conv2d(seg, state[prefix + "mlp_shared.0.weight"], state[prefix + "mlp_shared.0.bias"], padding=1))
    gamma = F.conv2d(actv, state[prefix + "mlp_gamma.weight"], state[prefix + "mlp_gamma.bias"], padding=1)
    beta = F.conv2d(actv, state[prefix + "mlp_beta.weight"], state[prefix + "mlp_beta.bias"], padding=1)
    return normalized * (1 + gamma) + beta


def norm_res_block(x: Tensor, seg: Tensor, state: State, prefix: str, training: bool, noise: NoiseSource) -> Tensor:
    """NormResBlock.forward with up_scale=False, architecture.py:343-357: the shortcut is the identity
    (norm_s / conv_s exist as parameters but are never executed)."""
    y = F.relu(spade(x, seg, state, prefix + "norm_0."))
    y = conv2d(y, conv_weight(state, prefix + "conv_0.", training), None, 1, 1, "reflect")
    y = noise_injection(y, state, prefix + "noise_0.", noise)
    y = F.relu(spade(y, seg, state, prefix + "norm_1."))
    y = conv2d(y, conv_weight(state, prefix + "conv_1.", training), None, 1, 1, "reflect")
    y = noise_injection(y, state, prefix + "noise_1.", noise)
    return y + x


def norm_conv_block(x: Tensor, seg: Tensor, state: State, prefix: str, training: bool, noise: NoiseSource) -> Tensor:
    """NormConvBlock.forward with up_scale=True, architecture.py:241-245:
    noise(conv(relu(SPADE(upsample2x_nearest(x)))))."""
    x = F.interpolate(x, scale_factor=2, mode="nearest")
    y = F.relu(spade(x, seg, state, prefix + "norm."))
    y = conv2d(y, conv_weight(state, prefix + "conv.", training), None, 1, 1, "reflect")
    return noise_injection(y, state, prefix + "noise.", noise)


# --------------------------------------------------------------------------------------------------
# networks
# --------------------------------------------------------------------------------------------------
def _count(state: State, stem: str) -> int:
    idx = set()
    for k in state:
        if k.startswith(stem):
            idx.add(int(k[len(stem):].split(".")[0]))
    return len(idx)


def generator_forward(state: State, x: Tensor, labels: Tensor, training: bool,
                      noise: Optional[NoiseSource] = None, cycle_gan: bool = False) -> Tuple[Tensor, Tensor]:
    """DefectGanGenerator.forward (skip_conn=False, spade), generator.py:243-275."""
    noise = noise if noise is not None else NoiseSource()
    seg = labels if labels.dim() == 4 else labels.reshape(labels.shape[0], labels.shape[1], 1, 1)
    feat = conv_block(x, state, "stem.conv_block.", 7, 1, 3, "reflect", True, "leaky_relu", training)
    for i in range(_count(state, "enc_blk.")):
        feat = conv_block(feat, state, f"enc_blk.{i}.conv_block.", 4, 2, 1, "reflect", True, "leaky_relu", training)
    for i in range(_count(state, "enc_res_blk.")):
        feat = res_block(feat, state, f"enc_res_blk.{i}.", training)
    for i in range(_count(state, "dec_res_blk.")):
        feat = norm_res_block(feat, seg, state, f"dec_res_blk.{i}.", training, noise)
    for i in range(_count(state, "dec_blk.")):
        feat = norm_conv_block(feat, seg, state, f"dec_blk.{i}.", training, noise)
    if feat.isnan().any():  # generator.py:266-267
        feat = feat.nan_to_num()
    fg = torch.tanh(conv2d(feat, state["foreground_head.de_conv_block.0.weight"], None, 1, 1, "reflect"))
    prob = torch.sigmoid(conv2d(feat, state["distribution_head.de_conv_block.0.weight"], None, 1, 1, "reflect"))
    out = x * (1 - prob) + fg * prob  # generator.py:270
    return (fg, prob) if cycle_gan else (out, prob)


def discriminator_forward(state: State, x: Tensor, training: bool) -> Tuple[Tensor, Tensor]:
    """DefectGanDiscriminator.forward, discriminator.py:92-98: (num_layers+1) x ConvBlock(4x4 s2 reflect,
    LReLU 0.2, spectral, no norm); src_clf 3x3 reflect 'same' -> 1; cls_clf (H/2^(L+1))-sized valid conv -> label_nc."""
    feat = x
    for i in range(_count(state, "enc_blk.")):
        feat = conv_block(feat, state, f"enc_blk.{i}.conv_block.", 4, 2, 1, "reflect", False, "leaky_relu", training)
    src = conv2d(feat, state["src_clf.conv_block.0.weight"], None, 1, 1, "reflect")
    cls = F.conv2d(feat, state["cls_clf.conv_block.0.weight"])
    return src, cls.reshape(cls.shape[0], cls.shape[1])


# --------------------------------------------------------------------------------------------------
# losses and steps
# --------------------------------------------------------------------------------------------------
def cal_loss(logits: Tensor, targets: Tensor, loss_type: str) -> Tensor:
    """BaseModel._cal_loss, models/base_model.py:68-80."""
    if loss_type in ("bce", "bce_logits"):
        return F.binary_cross_entropy_with_logits(logits, targets)
    if loss_type in ("cce", "cce_logits"):
        return F.cross_entropy(logits, targets)
    if loss_type == "l1":
        return F.l1_loss(logits, targets)
    if loss_type in ("l2", "mse"):
        return F.mse_loss(logits, targets)
    raise ValueError(loss_type)


def normal_labels(df_labels: Tensor) -> Tensor:
    """_get_label_and_style_feat, models/defectgan_model.py:413-415: one-hot class 0."""
    nm = torch.zeros_like(df_labels)
    nm[:, 0] = 1
    return nm


def discriminator_loss(g: State, d: State, bg: Tensor, df_labels: Tensor, df: Tensor,
                       clf_loss_type: str = "bce", noise: Optional[NoiseSource] = None) -> Tuple[Tensor, Tensor]:
    """DefectGanModel._compute_discriminator_loss, models/defectgan_model.py:251-292 (mode 'discriminator':
    netD.train(), netG.eval(), :87-90).  Returns (gan_loss, clf_loss)."""
    nm_labels = normal_labels(df_labels)
    with torch.no_grad():
        fake_df, _ = generator_forward(g, bg, df_labels, False, noise)
        fake_nm, _ = generator_forward(g, df, nm_labels, False, noise)
    fd_src, _ = discriminator_forward(d, fake_df, True)
    fn_src, _ = discriminator_forward(d, fake_nm, True)
    rd_src, rd_cls = discriminator_forward(d, df, True)
    rn_src, rn_cls = discriminator_forward(d, bg, True)
    ones, zeros = torch.ones_like(rd_src), torch.zeros_like(fd_src)
    gan = torch.stack([cal_loss(fd_src, zeros, "bce"), cal_loss(fn_src, zeros, "bce"),
                       cal_loss(rd_src, ones, "bce"), cal_loss(rn_src, ones, "bce")]).mean()
    clf = torch.stack([cal_loss(rd_cls, df_labels.view_as(rd_cls), clf_loss_type),
                       cal_loss(rn_cls, nm_labels.view_as(rn_cls), clf_loss_type)]).mean()
    return gan, clf


def generator_loss(g: State, d: State, bg: Tensor, df_labels: Tensor, df: Tensor,
                   clf_loss_type: str = "bce", noise: Optional[NoiseSource] = None):
    """DefectGanModel._compute_generator_loss, models/defectgan_model.py:173-249 (mode 'generator':
    netD.eval(), netG.train(), :83-86).  Returns (gan, clf, rec, sd_cyc, sd_con)."""
    nm_labels = normal_labels(df_labels)
    fake_df, df_prob = generator_forward(g, bg, df_labels, True, noise)
    rec_nm, rec_df_prob = generator_forward(g, fake_df, nm_labels, True, noise)
    fake_nm, nm_prob = generator_forward(g, df, nm_labels, True, noise)
    rec_df, rec_nm_prob = generator_forward(g, fake_nm, df_labels, True, noise)
    fd_src, fd_cls = discriminator_forward(d, fake_df, False)
    fn_src, fn_cls = discriminator_forward(d, fake_nm, False)
    ones = torch.ones_like(fd_src)
    gan = torch.stack([cal_loss(fd_src, ones, "bce"), cal_loss(fn_src, ones, "bce")]).mean()
    clf = torch.stack([cal_loss(fd_cls, df_labels.view_as(fd_cls), clf_loss_type),
                       cal_loss(fn_cls, nm_labels.view_as(fn_cls), clf_loss_type)]).mean()
    rec = torch.stack([cal_loss(rec_df, df, "l1"), cal_loss(rec_nm, bg, "l1")]).mean()
    cyc = torch.stack([cal_loss(df_prob, rec_df_prob, "l1"), cal_loss(nm_prob, rec_nm_prob, "l1")]).mean()
    z = torch.zeros_like(df_prob)
    con = torch.stack([cal_loss(df_prob, z, "l1"), cal_loss(nm_prob, z, "l1"),
                       cal_loss(rec_df_prob, z, "l1"), cal_loss(rec_nm_prob, z, "l1")]).mean()
    return gan, clf, rec, cyc, con


LOSS_WEIGHT_DEFAULT = (2, 5, 5, 5, 1)  # options/defectgan_options.py:100-101 [cls_d, cls_g, rec, sd_cyc, sd_con]


class Adam:
    """torch.optim.Adam(betas=(0.5, 0.999), eps=1e-8, weight_decay=0) restated
    (trainers/base_trainer.py:75-89); skips parameters whose grad is None, like torch."""

    def __init__(self, params: Sequence[Tensor], lr: float = 2e-4, betas=(0.5, 0.999), eps: float = 1e-8):
        self.params = list(params)
        self.lr, self.betas, self.eps = lr, betas, eps
        self.state = {}

    def zero_grad(self):
        for p in self.params:
            p.grad = None

    @torch.no_grad()
    def step(self):
        b1, b2 = self.betas
        for i, p in enumerate(self.params):
            if p.grad is None:
                continue
            st = self.state.setdefault(i, {"t": 0, "m": torch.zeros_like(p), "v": torch.zeros_like(p)})
            st["t"] += 1
            t = st["t"]
            st["m"].mul_(b1).add_(p.grad, alpha=1 - b1)
            st["v"].mul_(b2).addcmul_(p.grad, p.grad, value=1 - b2)
            denom = (st["v"].sqrt() / math.sqrt(1 - b2 ** t)).add_(self.eps)
            p.addcdiv_(st["m"], denom, value=-self.lr / (1 - b1 ** t))


def trainable_keys(state: State) -> List[str]:
    """Parameters (not buffers) in state-dict order = nn.Module.parameters() order."""
    skip = ("weight_u", "weight_v", "running_mean", "running_var", "num_batches_tracked")
    return [k for k in state if not k.endswith(skip)]


class Trainer:
    """DefectGanTrainer._train_discriminator_once / _train_generator_once restated
    (trainers/defectgan_trainer.py:170-180 / 138-168) over functional states."""

    def __init__(self, g: State, d: State, lr: float = 2e-4, loss_weight=LOSS_WEIGHT_DEFAULT,
                 clf_loss_type: str = "bce"):
        self.g, self.d = g, d
        for st in (g, d):
            for k in trainable_keys(st):
                st[k].requires_grad_(True)
        self.opt_g = Adam([g[k] for k in trainable_keys(g)], lr)
        self.opt_d = Adam([d[k] for k in trainable_keys(d)], lr)
        self.w = dict(zip(("clf_d", "clf_g", "rec", "sd_cyc", "sd_con"), loss_weight))
        self.clf_loss_type = clf_loss_type

    def train_discriminator_once(self, bg, df_labels, df, noise=None):
        self.opt_d.zero_grad()
        gan, clf = discriminator_loss(self.g, self.d, bg, df_labels, df, self.clf_loss_type, noise)
        (gan + clf * self.w["clf_d"]).backward()
        self.opt_d.step()
        return gan.item(), clf.item()

    def train_generator_once(self, bg, df_labels, df, noise=None):
        self.opt_g.zero_grad()
        gan, clf, rec, cyc, con = generator_loss(self.g, self.d, bg, df_labels, df, self.clf_loss_type, noise)
        (gan + clf * self.w["clf_g"] + rec * self.w["rec"] + cyc * self.w["sd_cyc"] + con * self.w["sd_con"]).backward()
        self.opt_g.step()
        return gan.item(), clf.item(), rec.item(), cyc.item(), con.item()


# --------------------------------------------------------------------------------------------------
# synthetic inputs / random-init states (no reference needed: travels to the GPU box)
# --------------------------------------------------------------------------------------------------
def synthetic_batch(batch: int, image_size: int, label_nc: int = 6, seed: int = 0):
    """SURVEY.md section 8(d) synthetic inputs: bg, df = randn(B,3,H,H).clamp(-1,1); multi-hot defect labels with
    column 0 forced to 0."""
    gen = torch.Generator().manual_seed(seed)
    bg = torch.randn(batch, 3, image_size, image_size, generator=gen).clamp_(-1, 1)
    df = torch.randn(batch, 3, image_size, image_size, generator=gen).clamp_(-1, 1)
    labels = (torch.rand(batch, label_nc, generator=gen) > 0.5).float()
    labels[:, 0] = 0
    return bg, labels, df


def _sn(state: State, prefix: str, shape, gen, std):
    w = torch.randn(*shape, generator=gen) * std
    state[prefix + "weight_orig"] = w
    state[prefix + "weight_u"] = _normalize(torch.randn(shape[0], generator=gen), 1e-12)
    state[prefix + "weight_v"] = _normalize(torch.randn(w[0].numel(), generator=gen), 1e-12)


def _plain(state: State, prefix: str, shape, gen, std, bias=False):
    state[prefix + "weight"] = torch.randn(*shape, generator=gen) * std
    if bias:
        state[prefix + "bias"] = torch.zeros(shape[0])


def _bn(state: State, prefix: str, c, gen, std):
    state[prefix + "weight"] = 1.0 + torch.randn(c, generator=gen) * std
    state[prefix + "bias"] = torch.zeros(c)
    state[prefix + "running_mean"] = torch.zeros(c)
    state[prefix + "running_var"] = torch.ones(c)
    state[prefix + "num_batches_tracked"] = torch.zeros((), dtype=torch.long)


def _spade(state: State, prefix: str, label_nc, c, hidden, gen, std):
    _plain(state, prefix + "mlp_shared.0.", (hidden, label_nc, 3, 3), gen, std, bias=True)
    _plain(state, prefix + "mlp_gamma.", (c, hidden, 3, 3), gen, std, bias=True)
    _plain(state, prefix + "mlp_beta.", (c, hidden, 3, 3), gen, std, bias=True)


def random_generator_state(ngf=64, num_scales=2, num_res=6, label_nc=6, hidden_nc=128, use_spectral=True,
                           add_noise=True, seed=1, std=0.02, noise_weight=0.0) -> State:
    """A state dict with the key set / shapes / order of DefectGanGenerator(opt).state_dict()
    (generator.py:52-241) filled with N(0, std) values (BaseNetwork.init_weights, base_network.py:27-61).
    The RNG stream differs from the reference's; use it only where the oracle itself is the checker."""
    gen = torch.Generator().manual_seed(seed)
    s: State = {}
    conv = _sn if use_spectral else (lambda st, p, shape, g, sd: _plain(st, p, shape, g, sd))
    conv(s, "stem.conv_block.0.", (ngf, 3, 7, 7), gen, std)
    _bn(s, "stem.conv_block.1.", ngf, gen, std)
    c = ngf
    for i in range(num_scales):
        conv(s, f"enc_blk.{i}.conv_block.0.", (2 * c, c, 4, 4), gen, std)
        _bn(s, f"enc_blk.{i}.conv_block.1.", 2 * c, gen, std)
        c *= 2
    for i in range(num_res // 2):
        for j in range(2):
            conv(s, f"enc_res_blk.{i}.res_block.{j}.conv_block.0.", (c, c, 3, 3), gen, std)
            _bn(s, f"enc_res_blk.{i}.res_block.{j}.conv_block.1.", c, gen, std)
    for i in range(num_res // 2):
        p = f"dec_res_blk.{i}."
        if add_noise:
            s[p + "noise_0.weight"] = torch.full((1, 1, 1, 1), float(noise_weight))
            s[p + "noise_1.weight"] = torch.full((1, 1, 1, 1), float(noise_weight))
        for n in ("norm_0.", "norm_1.", "norm_s."):
            _spade(s, p + n, label_nc, c, hidden_nc, gen, std)
        for n in ("conv_0.", "conv_1.", "conv_s."):
            conv(s, p + n, (c, c, 3, 3), gen, std)
    for i in range(num_scales):
        p = f"dec_blk.{i}."
        if add_noise:
            s[p + "noise.weight"] = torch.full((1, 1, 1, 1), float(noise_weight))
        _spade(s, p + "norm.", label_nc, c, hidden_nc, gen, std)
        conv(s, p + "conv.", (c // 2, c, 3, 3), gen, std)
        c //= 2
    _plain(s, "foreground_head.de_conv_block.0.", (3, c, 3, 3), gen, std)
    _plain(s, "distribution_head.de_conv_block.0.", (1, c, 3, 3), gen, std)
    return s


def random_discriminator_state(ndf=64, num_layers=5, label_nc=6, image_size=128, use_spectral=True,
                               seed=2, std=0.02) -> State:
    """Key set / shapes of DefectGanDiscriminator(opt).state_dict() (discriminator.py:49-90)."""
    gen = torch.Generator().manual_seed(seed)
    s: State = {}
    conv = _sn if use_spectral else (lambda st, p, shape, g, sd: _plain(st, p, shape, g, sd))
    conv(s, "enc_blk.0.conv_block.0.", (ndf, 3, 4, 4), gen, std)
    c = ndf
    for i in range(num_layers):
        conv(s, f"enc_blk.{i + 1}.conv_block.0.", (2 * c, c, 4, 4), gen, std)
        c *= 2
    k = image_size // (2 ** (num_layers + 1))
    _plain(s, "cls_clf.conv_block.0.", (label_nc, c, k, k), gen, std)
    _plain(s, "src_clf.conv_block.0.", (1, c, 3, 3), gen, std)
    return s


def clone_state(state: State) -> State:
    return {k: v.detach().clone() for k, v in state.items()}
