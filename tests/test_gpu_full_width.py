"""Parity at the BASELINE configurations, FULL width (ngf = ndf = 64, hidden_nc = 128) - the kernels that carry the
benchmarked step (conv_tma<256,4,1> / <128,4,2> / <64,5,2>, wgrad_tma<2,3> / <1,4>, the row-packed stem, out_pad SPADE,
BatchNorm-written padding) inside whole networks and at the bench's own tensor sizes.

  * config 1 (BASELINE.json configs[0]: 128x128, batch 4, --add_noise --use_spectral): generator forward + backward,
    discriminator forward + backward and one D-step + G-step through DefectGanModel + FlatAdam against
    oracle/defectgan_ref.py run in FP32 on the box's host cores (same state dict, inputs, injected noise);
  * config 2 (configs[1]: 256x256, batch 32): the discriminator loss (2 G eval passes + 4 D passes, forward + backward) and the
    generator loss forward at the full batch, and a full G-step backward at batch 8, against the same oracle moved to the
    GPU (stock torch FP32, TF32 disabled - a test checker, never on the product path);
  * every convolution geometry of the bench's per-geometry table (N = 32) forward / dgrad / wgrad, the SPADE class-map,
    BatchNorm and reflect-fold kernels at 32 x 258 x 258 x 128 against torch FP32 on the GPU, with an assert on WHICH kernel
    variant ran (dei2i_last_kernel).

Tolerances as tests/test_gpu_golden.py / test_gpu_kernels.py: 2e-2 (BF16 activations, losses), gradients by _grad_check.
"""
import json
import os

import pytest
import torch
import torch.nn.functional as F

from conftest import rel_err, ROOT
from test_gpu_golden import _grad_check

pytestmark = pytest.mark.gpu

TOL_ACT = 2e-2
TOL_BF16 = 1e-2
TOL_F32 = 2e-3


def _report(name, **kv):
    try:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "full_width_parity.jsonl"), "a") as f:
            f.write(json.dumps(dict(test=name, **kv)) + "\n")
    except OSError:
        pass


@pytest.fixture(autouse=True)
def _exact_fp32_checker():
    """The torch-CUDA checker computes in true FP32 (no TF32)."""
    old = (torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32)
    torch.backends.cudnn.allow_tf32 = False
    torch.backends.cuda.matmul.allow_tf32 = False
    yield
    torch.backends.cudnn.allow_tf32, torch.backends.cuda.matmul.allow_tf32 = old


def _opt(image_size, batch):
    from de_i2i_gan_b200 import options
    opt = options.readme_recipe(image_size=image_size, batch_size=batch, num_critics=1)
    opt.device = torch.device("cuda", 0)
    return opt


def _noise_shapes(b, h, passes):
    one = [(b, 1, h // 4, h // 4)] * 6 + [(b, 1, h // 2, h // 2), (b, 1, h, h)]
    return one * passes


def _draw(shapes, seed):
    g = torch.Generator().manual_seed(seed)
    return [torch.randn(s, generator=g) for s in shapes]


def _inject(noises):
    from de_i2i_gan_b200.models.networks import architecture as A
    q = [n.clone() for n in noises]

    def src(n, h, w, device):
        t = q.pop(0)
        assert tuple(t.shape) == (n, 1, h, w), (t.shape, (n, 1, h, w))
        return t
    A.set_noise_source(src)
    return q


def _clear():
    from de_i2i_gan_b200.models.networks import architecture as A
    A.set_noise_source(None)


def _states(image_size, noise_weight=0.1, sn_warm=5):
    """Random-init full-width states.  The random u / v of the spectral-norm buffers are advanced by a few power iterations
    first: with un-iterated u, v sigma is underestimated by orders of magnitude, the six stacked discriminator layers then
    produce logits of 1e10 and every comparison downstream is about an ill-conditioned graph rather than about the kernels."""
    from oracle import defectgan_ref as R
    g, d = R.random_generator_state(noise_weight=noise_weight), R.random_discriminator_state(image_size=image_size)
    for st in (g, d):
        for k in list(st):
            if k.endswith("weight_orig"):
                for _ in range(sn_warm):
                    R.spectral_weight(st, k[: -len("weight_orig")], True)
    return g, d


def _to(state, dev):
    return {k: v.detach().clone().to(dev) for k, v in state.items()}


def _req_grad(state):
    from oracle import defectgan_ref as R
    for k in R.trainable_keys(state):
        state[k].requires_grad_(True)
    return state


def _ref_grads(state):
    from oracle import defectgan_ref as R
    return {k: state[k].grad.detach().float().cpu() for k in R.trainable_keys(state) if state[k].grad is not None}


class _KernelLog:
    def __enter__(self):
        from de_i2i_gan_b200 import ops
        ops.KERNEL_LOG = self.log = []
        return self

    def __exit__(self, *exc):
        from de_i2i_gan_b200 import ops
        ops.KERNEL_LOG = None
        return False

    def variants(self):
        return sorted({k for _, _, names in self.log for k in names.split("+") if k})


# ======================================================================================================== config 1
def test_c1_generator_forward_backward_vs_cpu_oracle():
    """defectGAN/models/networks/generator.py:243-275 at full width, 128x128, batch 4."""
    from oracle import defectgan_ref as R
    from de_i2i_gan_b200.models.networks.generator import DefectGanGenerator
    from de_i2i_gan_b200.synthetic import synthetic_batch
    g_state, _ = _states(128)
    bg, labels, _ = synthetic_batch(4, 128, seed=0)
    noises = _draw(_noise_shapes(4, 128, 1), 7)
    G = DefectGanGenerator(_opt(128, 4))
    G.load_state_dict(g_state)
    G.cuda().train()
    q = _inject(noises)
    try:
        with _KernelLog() as kl:
            out, prob = G(bg.cuda(), labels.cuda())
            (out.square().mean() + prob.mean()).backward()
        assert not q
    finally:
        _clear()
    st = _req_grad(R.clone_state(g_state))
    o_ref, p_ref = R.generator_forward(st, bg, labels, True, R.NoiseSource(noises))
    (o_ref.square().mean() + p_ref.mean()).backward()
    e_out, e_prob = rel_err(out, o_ref), rel_err(prob, p_ref)
    worst = _grad_check(G.named_parameters(), _ref_grads(st), "C1 G")
    bufs = G.state_dict()
    buf_keys = [k for k in st if k.endswith(("weight_u", "weight_v", "running_mean", "running_var"))]
    e_buf = max(rel_err(bufs[k], st[k]) for k in buf_keys)
    used = kl.variants()
    _report("c1/generator", out=e_out, prob=e_prob, grads=worst, buffers=e_buf, kernels=used)
    assert e_out < TOL_ACT and e_prob < TOL_ACT, (e_out, e_prob)
    assert e_buf < 1e-2, e_buf
    # the production tcgen05 / TMA variants, not the narrow-channel fallbacks, carried this network
    assert any(k.startswith("conv_tma_pair<256") for k in used) and any(k.startswith("wgrad_tma<2") for k in used), used
    assert any(k.startswith("wgrad_tma<1") for k in used), used


def test_c1_discriminator_forward_backward_vs_cpu_oracle():
    """defectGAN/models/networks/discriminator.py:92-98 at full width, 128x128, batch 4."""
    from oracle import defectgan_ref as R
    from de_i2i_gan_b200.models.networks.discriminator import DefectGanDiscriminator
    from de_i2i_gan_b200.synthetic import synthetic_batch
    _, d_state = _states(128)
    x, _, _ = synthetic_batch(4, 128, seed=1)
    D = DefectGanDiscriminator(_opt(128, 4))
    D.load_state_dict(d_state)
    D.cuda().train()
    with _KernelLog() as kl:
        src, cls = D(x.cuda())
        (src.mean() + cls.square().mean()).backward()
    sd = _req_grad(R.clone_state(d_state))
    s_ref, c_ref = R.discriminator_forward(sd, x, True)
    (s_ref.mean() + c_ref.square().mean()).backward()
    e_src, e_cls = rel_err(src, s_ref), rel_err(cls, c_ref)
    worst = _grad_check(D.named_parameters(), _ref_grads(sd), "C1 D")
    _report("c1/discriminator", src=e_src, cls=e_cls, grads=worst, kernels=kl.variants())
    assert e_src < TOL_ACT and e_cls < TOL_ACT, (e_src, e_cls)


def test_c1_d_step_and_g_step_vs_cpu_oracle():
    """defectGAN/trainers/defectgan_trainer.py:138-180 on BASELINE config 1: one D-step + one G-step through DefectGanModel +
    FlatAdam vs the oracle's Trainer (torch-Adam restated) on the host cores: all 7 losses, the G-step gradients, and the
    direction of both parameter updates.  Adam's first update is -lr * sign(g) element-wise (m / sqrt(v) = +-1), so the cosine of
    the two update vectors is 1 - 2 * (fraction of parameters whose gradient sign differs): elements whose gradient is below the
    BF16 noise floor flip (measured: 8 % of G's 16 M, 4 % of D's 45 M parameters) while the gradient itself agrees to 2e-2
    relative L2 (checked by _grad_check) - the bar is therefore 0.8, not the 0.9 of the reduced-width golden test."""
    from oracle import defectgan_ref as R
    from de_i2i_gan_b200.models.defectgan_model import DefectGanModel
    from de_i2i_gan_b200.optim import FlatAdam
    from de_i2i_gan_b200.synthetic import synthetic_batch
    g_state, d_state = _states(128)
    bg, labels, df = synthetic_batch(4, 128, seed=0)
    noises_d = _draw(_noise_shapes(4, 128, 2), 11)
    noises_g = _draw(_noise_shapes(4, 128, 4), 12)
    model = DefectGanModel(_opt(128, 4))
    model.netG.load_state_dict(g_state)
    model.netD.load_state_dict(d_state)
    model.netG.cuda()
    model.netD.cuda()
    optD = FlatAdam(list(model.netD.parameters()), lr=2e-4, betas=(0.5, 0.999))
    optG = FlatAdam(list(model.netG.parameters()), lr=2e-4, betas=(0.5, 0.999))
    w = dict(zip(("clf_d", "clf_g", "rec", "sd_cyc", "sd_con"), R.LOSS_WEIGHT_DEFAULT))
    b, l, d = bg.cuda(), labels.cuda(), df.cuda()
    try:
        q = _inject(noises_d)
        optD.zero_grad()
        gan_d, clf_d = model("discriminator", b, l, d)
        (gan_d + clf_d * w["clf_d"]).backward()
        optD.step()
        assert not q
        q = _inject(noises_g)
        optG.zero_grad()
        gan, clf, rec, cyc, con = model("generator", b, l, d)
        (gan + clf * w["clf_g"] + rec * w["rec"] + cyc * w["sd_cyc"] + con * w["sd_con"]).backward()
        assert not q
    finally:
        _clear()
    got = dict(gan_d=gan_d.item(), clf_d=clf_d.item(), gan_g=gan.item(), clf_g=clf.item(), rec=rec.item(), cyc=cyc.item(), con=con.item())
    # oracle
    g_ref, d_ref = R.clone_state(g_state), R.clone_state(d_state)
    tr = R.Trainer(g_ref, d_ref)
    r = {}
    r["gan_d"], r["clf_d"] = tr.train_discriminator_once(bg, labels, df, R.NoiseSource(noises_d))
    tr.opt_g.zero_grad()
    losses = R.generator_loss(tr.g, tr.d, bg, labels, df, "bce", R.NoiseSource(noises_g))
    (losses[0] + losses[1] * w["clf_g"] + losses[2] * w["rec"] + losses[3] * w["sd_cyc"] + losses[4] * w["sd_con"]).backward()
    r.update(dict(zip(("gan_g", "clf_g", "rec", "cyc", "con"), [x.item() for x in losses])))
    worst = _grad_check(model.netG.named_parameters(), _ref_grads(tr.g), "C1 G-step")
    optG.step()
    tr.opt_g.step()
    errs = {k: abs(got[k] - r[k]) / max(abs(r[k]), 1e-3) for k in r}
    cos = {}
    for name, net, before, after in (("g", model.netG, g_state, tr.g), ("d", model.netD, d_state, tr.d)):
        keys = [n for n, _ in net.named_parameters()]
        st = net.state_dict()
        d_r = torch.cat([(after[k].detach() - before[k]).flatten() for k in keys])
        d_g = torch.cat([(st[k].detach().cpu() - before[k]).flatten() for k in keys])
        cos[name] = (torch.dot(d_r, d_g) / (d_r.norm() * d_g.norm())).item()
    _report("c1/d_step_g_step", losses=errs, grads=worst, cos_g=cos["g"], cos_d=cos["d"])
    assert max(errs.values()) < TOL_ACT, (errs, got, r)
    assert cos["g"] > 0.8 and cos["d"] > 0.8, cos


# ======================================================================================================== config 2
def _c2_model(batch):
    from de_i2i_gan_b200.models.defectgan_model import DefectGanModel
    g_state, d_state = _states(256)
    model = DefectGanModel(_opt(256, batch))
    model.netG.load_state_dict(g_state)
    model.netD.load_state_dict(d_state)
    model.netG.cuda()
    model.netD.cuda()
    return model, g_state, d_state


def test_c2_discriminator_loss_full_batch_vs_fp32_torch():
    """BASELINE config 2 (256x256, batch 32): _compute_discriminator_loss (defectgan_model.py:251-292) forward + backward
    at the bench's own tensor sizes (32 x 258 x 258 x 128 BF16 = 545 MB tensors, byte offsets > 2^31, MS = 2 tiles, out_pad
    SPADE at 258^2, row-packed stem at 256^2) vs the oracle on the GPU in FP32."""
    from oracle import defectgan_ref as R
    from de_i2i_gan_b200.synthetic import synthetic_batch
    B = 32
    model, g_state, d_state = _c2_model(B)
    bg, labels, df = synthetic_batch(B, 256, seed=0)
    noises = _draw(_noise_shapes(B, 256, 2), 21)
    b, l, d = bg.cuda(), labels.cuda(), df.cuda()
    q = _inject(noises)
    try:
        with _KernelLog() as kl:
            gan, clf = model("discriminator", b, l, d)
            (gan + 2 * clf).backward()
        assert not q
    finally:
        _clear()
    dev = torch.device("cuda", 0)
    g_ref, d_ref = _to(g_state, dev), _req_grad(_to(d_state, dev))
    gan_r, clf_r = R.discriminator_loss(g_ref, d_ref, b, l, d, "bce", R.NoiseSource([n.cuda() for n in noises]))
    (gan_r + 2 * clf_r).backward()
    e = dict(gan=abs(gan.item() - gan_r.item()) / abs(gan_r.item()), clf=abs(clf.item() - clf_r.item()) / abs(clf_r.item()))
    worst = _grad_check(model.netD.named_parameters(), _ref_grads(d_ref), "C2 D-loss")
    used = kl.variants()
    _report("c2/discriminator_loss_b32", losses=e, grads=worst, kernels=used)
    assert max(e.values()) < TOL_ACT, e
    for want in ("conv_tma_pair<256,6,1>", "conv_tma<128,4,2>", "conv_tma<64,5,2>"):
        assert want in used, (want, used)


def test_c2_generator_loss_forward_full_batch_vs_fp32_torch():
    """_compute_generator_loss (defectgan_model.py:173-249) forward at 256x256, batch 32 (4 G passes in train mode + 2 D passes):
    all five loss terms and the four generator outputs vs the oracle on the GPU in FP32."""
    from oracle import defectgan_ref as R
    from de_i2i_gan_b200.synthetic import synthetic_batch
    B = 32
    model, g_state, d_state = _c2_model(B)
    bg, labels, df = synthetic_batch(B, 256, seed=1)
    noises = _draw(_noise_shapes(B, 256, 4), 22)
    b, l, d = bg.cuda(), labels.cuda(), df.cuda()
    q = _inject(noises)
    try:
        with torch.no_grad():
            got = [x.item() for x in model("generator", b, l, d)]
        assert not q
    finally:
        _clear()
    dev = torch.device("cuda", 0)
    with torch.no_grad():
        ref = [x.item() for x in R.generator_loss(_to(g_state, dev), _to(d_state, dev), b, l, d, "bce",
                                                  R.NoiseSource([n.cuda() for n in noises]))]
    errs = {k: abs(a - r) / max(abs(r), 1e-3) for k, a, r in zip(("gan", "clf", "rec", "cyc", "con"), got, ref)}
    _report("c2/generator_loss_fwd_b32", losses=errs, got=got, ref=ref)
    assert max(errs.values()) < TOL_ACT, (errs, got, ref)


def test_c2_generator_step_backward_batch8_vs_fp32_torch():
    """The G-step loss graph at 256x256 with gradients (batch 8: the FP32 torch autograd graph of four generator passes at
    batch 32 does not fit beside ours): whole-network gradient check of the generator."""
    from oracle import defectgan_ref as R
    from de_i2i_gan_b200.synthetic import synthetic_batch
    B = 8
    model, g_state, d_state = _c2_model(B)
    bg, labels, df = synthetic_batch(B, 256, seed=2)
    noises = _draw(_noise_shapes(B, 256, 4), 23)
    b, l, d = bg.cuda(), labels.cuda(), df.cuda()
    w = dict(zip(("clf_d", "clf_g", "rec", "sd_cyc", "sd_con"), R.LOSS_WEIGHT_DEFAULT))
    for p in model.netD.parameters():
        p.requires_grad_(False)
    q = _inject(noises)
    try:
        gan, clf, rec, cyc, con = model("generator", b, l, d)
        (gan + clf * w["clf_g"] + rec * w["rec"] + cyc * w["sd_cyc"] + con * w["sd_con"]).backward()
        assert not q
    finally:
        _clear()
    got = [x.item() for x in (gan, clf, rec, cyc, con)]
    dev = torch.device("cuda", 0)
    g_ref, d_ref = _req_grad(_to(g_state, dev)), _to(d_state, dev)
    lr = R.generator_loss(g_ref, d_ref, b, l, d, "bce", R.NoiseSource([n.cuda() for n in noises]))
    (lr[0] + lr[1] * w["clf_g"] + lr[2] * w["rec"] + lr[3] * w["sd_cyc"] + lr[4] * w["sd_con"]).backward()
    ref = [x.item() for x in lr]
    errs = {k: abs(a - r) / max(abs(r), 1e-3) for k, a, r in zip(("gan", "clf", "rec", "cyc", "con"), got, ref)}
    worst = _grad_check(model.netG.named_parameters(), _ref_grads(g_ref), "C2 G-step b8")
    _report("c2/generator_step_b8", losses=errs, grads=worst)
    assert max(errs.values()) < TOL_ACT, errs


# ======================================================================================================== C2 geometry sweep
def _rel_dev(a, b):
    """max |a-b| / max |b| on the device (the tensors are up to 1 GB)."""
    return ((a.float() - b.float()).abs().max() / b.float().abs().max().clamp_min(1e-20)).item()


def _nhwc(x_nchw, c_ld=None):
    """fp32 NCHW CUDA -> bf16 NHWC CUDA with the channel pitch padded to 8 (plain torch, independent of our kernels)."""
    n, c, h, w = x_nchw.shape
    c_ld = c_ld or (c + 7) // 8 * 8
    y = torch.zeros(n, h, w, c_ld, dtype=torch.bfloat16, device=x_nchw.device)
    y[..., :c] = x_nchw.permute(0, 2, 3, 1)
    return y


def _nchw(y_nhwc, c):
    return y_nhwc[..., :c].permute(0, 3, 1, 2).float()


# name: (N, Cin, H, W, Cout, kh, kw, stride, expected fwd / dgrad / wgrad kernel variant (prefix match; None = just record))
# Every row of profiles/r01m_conv_breakdown_head.json (the bench's per-geometry table at 256x256, batch 32); inputs carry their
# reflect padding already ('valid' convolutions), exactly as the networks run them.
C2_GEOMETRIES = {
    "res_256_256_3x3_in66": (32, 256, 66, 66, 256, 3, 3, 1, "conv_tma_pair<256,6,1>", "conv_tma_pair<256,6,1>", "wgrad_tma<2,3>"),
    "dec1_128_64_3x3_in258": (32, 128, 258, 258, 64, 3, 3, 1, "conv_tma_pair<64,5,2>", "conv_tma_pair<128,5,2>", "wgrad_tma<1,4>"),
    "dec0_256_128_3x3_in130": (32, 256, 130, 130, 128, 3, 3, 1, "conv_tma_pair<128,5,2>", "conv_tma_pair<256,6,1>", "wgrad_tma<1,4>"),
    "enc0_64_128_4x4s2_in258": (32, 64, 258, 258, 128, 4, 4, 2, "conv_tma_pair<128,5,2>", "conv_tma<64", "wgrad_tma<1,4>"),
    "enc1_128_256_4x4s2_in130": (32, 128, 130, 130, 256, 4, 4, 2, "conv_tma_pair<256,6,1>", "conv_tma<128", "wgrad_tma<2,3>"),
    "d1_64_128_4x4s2_in130": (32, 64, 130, 130, 128, 4, 4, 2, "conv_tma<128", "conv_tma<64", "wgrad_tma<1,4>"),
    "d2_128_256_4x4s2_in66": (32, 128, 66, 66, 256, 4, 4, 2, "conv_tma_pair<256,6,1>", "conv_tma<128", "wgrad_tma<2,3>"),
    "d3_256_512_4x4s2_in34": (32, 256, 34, 34, 512, 4, 4, 2, "conv_tma_pair<", "conv_tma", "wgrad_tma<2,3>"),
    "d4_512_1024_4x4s2_in18": (32, 512, 18, 18, 1024, 4, 4, 2, "conv_tma_pair<", "conv_tma", "wgrad_tma<2,3>"),
    "d5_1024_2048_4x4s2_in10": (32, 1024, 10, 10, 2048, 4, 4, 2, "conv_tma_pair<256,6,1>", "conv_tma", "wgrad_tma<2,3>"),
    "stem_rowpacked_64_64_7x1_in262x256": (32, 64, 262, 256, 64, 7, 1, 1, "conv_ring<7>", "conv_ring<7>", "wgrad_ring<7>"),
    "heads_64_4_3x3_in258": (32, 64, 258, 258, 4, 3, 3, 1, "conv_heads_fwd", "conv_c8_fwd", "wgrad_heads_mma"),
    "d0_3_64_4x4s2_in258": (32, 3, 258, 258, 64, 4, 4, 2, "conv_c8_fwd", None, "wgrad_c8"),
    "spade_gb_128_512_3x3_in5": (32, 128, 5, 5, 512, 3, 3, 1, None, None, None),
    "d_src_2048_1_3x3_in6": (32, 2048, 6, 6, 1, 3, 3, 1, "conv_narrow_fwd", None, None),
    "d_cls_2048_6_4x4_in4": (32, 2048, 4, 4, 6, 4, 4, 1, "conv_narrow_fwd", None, None),
}


@pytest.mark.parametrize("name", list(C2_GEOMETRIES))
def test_c2_geometry_fwd_dgrad_wgrad(name):
    """One convolution of the 256x256 / batch-32 step, forward + dgrad + wgrad through the C ABI, vs torch FP32 (cuDNN, TF32 off)
    on the same BF16-rounded operands; asserts the kernel variant that ran."""
    from de_i2i_gan_b200 import ops
    n, cin, h, w, cout, kh, kw, stride, k_fwd, k_dg, k_wg = C2_GEOMETRIES[name]
    pad = 1 if name.startswith("spade_gb") else 0
    g = torch.Generator(device="cuda").manual_seed(5)
    x = torch.randn(n, cin, h, w, generator=g, device="cuda").bfloat16().float()
    wt = (torch.randn(cout, cin, kh, kw, generator=g, device="cuda") / (cin * kh * kw) ** 0.5).bfloat16().float()
    x.requires_grad_(True)
    wt.requires_grad_(True)
    y_ref = F.conv2d(x, wt, None, stride=stride, padding=pad)
    dy = torch.randn(y_ref.shape, generator=g, device="cuda").bfloat16().float()
    dx_ref, dw_ref = torch.autograd.grad(y_ref, (x, wt), dy)
    y_ref = y_ref.detach()
    spec = ops.ConvSpec(cin, cout, kh, stride, pad, "zeros", kw if kw != kh else 0)
    w_fwd, w_dg = ops.pack_weight(wt.detach().contiguous(), spec)
    xd, dyd = _nhwc(x.detach()), _nhwc(dy)
    errs, kern = {}, {}
    y = ops.conv2d_fwd(xd, w_fwd, spec)
    kern["fwd"] = ops.last_kernel()
    errs["fwd"] = _rel_dev(_nchw(y, cout), y_ref)
    dx = ops.conv2d_dgrad(dyd, w_dg, spec, (h, w))
    kern["dgrad"] = ops.last_kernel()
    errs["dgrad"] = _rel_dev(_nchw(dx, cin), dx_ref)
    dw = torch.empty(cout, cin, kh, kw, device="cuda")
    ops.conv2d_wgrad(xd, dyd, dw, spec)
    kern["wgrad"] = ops.last_kernel()
    errs["wgrad"] = _rel_dev(dw, dw_ref)
    torch.cuda.synchronize()
    _report("c2_geometry/" + name, **errs, kernels=kern)
    assert errs["fwd"] < TOL_BF16 and errs["dgrad"] < TOL_BF16 and errs["wgrad"] < TOL_F32, (errs, kern)
    for op, want in (("fwd", k_fwd), ("dgrad", k_dg), ("wgrad", k_wg)):
        if want is not None:
            assert kern[op].startswith(want), (op, kern[op], want)


def _cls_index(n, dev):
    return torch.tensor([i if i < 2 else (i - n + 5 if i >= n - 2 else 2) for i in range(n)], device=dev)


@pytest.mark.parametrize("c,hx,up,n", [(128, 128, 2, 32), (256, 64, 2, 32), (256, 64, 1, 32)])
def test_c2_spade_cls_out_pad_fwd_bwd(c, hx, up, n):
    """dei2i_spade_cls_fwd / _bwd at the bench's sizes (dec_blk.1: x 32x128x128x128 -> out 32x258x258x128 with the consumer's
    reflect padding; dec_blk.0; dec_res_blk) vs torch FP32 on the GPU: IN(up2x(x)) * (1 + gamma) + beta, ReLU, reflect pad."""
    from de_i2i_gan_b200 import ops
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cuda").manual_seed(16)
    h = hx * up
    x = (torch.randn(n, c, hx, hx, generator=g, device=dev) * 2 + 1).bfloat16().float().requires_grad_(True)
    gamma5 = (torch.randn(n, c, 5, 5, generator=g, device=dev) * 0.5).bfloat16().float().requires_grad_(True)
    beta5 = (torch.randn(n, c, 5, 5, generator=g, device=dev) * 0.5).bfloat16().float().requires_grad_(True)
    idx = _cls_index(h, dev)
    xu = F.interpolate(x, scale_factor=up, mode="nearest") if up > 1 else x
    out_ref = F.relu(F.instance_norm(xu, eps=1e-5) * (1 + gamma5[:, :, idx][:, :, :, idx]) + beta5[:, :, idx][:, :, :, idx])
    out_ref = F.pad(out_ref, (1, 1, 1, 1), mode="reflect")
    dout = torch.randn(out_ref.shape, generator=g, device=dev).bfloat16().float()
    dx_ref, dg_ref, db_ref = torch.autograd.grad(out_ref, (x, gamma5, beta5), dout)
    out_ref = out_ref.detach()
    xd, gbd = _nhwc(x.detach()), _nhwc(torch.cat([gamma5.detach(), beta5.detach()], 1))
    out, mr = ops.spade_cls_fwd(xd, gbd, up, out_pad=1)
    e_out = _rel_dev(_nchw(out, c), out_ref)
    dgb5, dx = ops.spade_cls_bwd(xd, gbd, out, _nhwc(dout), mr, up, out_pad=1)
    torch.cuda.synchronize()
    dgb = dgb5.permute(0, 3, 1, 2)
    e = dict(out=e_out, dx=_rel_dev(_nchw(dx, c), dx_ref), dgamma=_rel_dev(dgb[:, :c], dg_ref), dbeta=_rel_dev(dgb[:, c:], db_ref))
    _report(f"c2_spade_cls/c{c}_hx{hx}_up{up}", **e)
    assert e["out"] < TOL_BF16 and e["dx"] < 2e-2 and e["dgamma"] < 5e-3 and e["dbeta"] < 5e-3, e


@pytest.mark.parametrize("c,h,pad,act", [(64, 256, 1, "leaky_relu"), (128, 128, 1, "leaky_relu"), (256, 64, 1, None)])
def test_c2_batchnorm_pad_fwd_bwd(c, h, pad, act):
    """dei2i_bn_fwd_pad + dei2i_reflect_fold + dei2i_bn_bwd at the bench's sizes (stem 32x256x256x64, enc 128 @128^2, 256 @64^2)
    vs torch FP32 batch_norm (+ LeakyReLU, + residual for the act=None case) + reflect pad."""
    from de_i2i_gan_b200 import ops
    dev = torch.device("cuda", 0)
    n = 32
    g = torch.Generator(device="cuda").manual_seed(4)
    x = (torch.randn(n, c, h, h, generator=g, device=dev) * 1.5 + 0.3).bfloat16().float().requires_grad_(True)
    gamma = (1 + 0.1 * torch.randn(c, generator=g, device=dev)).requires_grad_(True)
    beta = (0.1 * torch.randn(c, generator=g, device=dev)).requires_grad_(True)
    res = torch.randn(n, c, h, h, generator=g, device=dev).bfloat16().float() if act is None else None
    rm, rv = torch.zeros(c, device=dev), torch.ones(c, device=dev)
    rm_ref, rv_ref = rm.clone(), rv.clone()
    y_ref = F.batch_norm(x, rm_ref, rv_ref, gamma, beta, True, 0.1, 1e-5)
    if act:
        y_ref = F.leaky_relu(y_ref, 0.2)
    if res is not None:
        y_ref = y_ref + res
    y_ref = F.pad(y_ref, (pad,) * 4, mode="reflect")
    dy = torch.randn(y_ref.shape, generator=g, device=dev).bfloat16().float()
    dx_ref, dg_ref, db_ref = torch.autograd.grad(y_ref, (x, gamma, beta), dy)
    xd = _nhwc(x.detach())
    y, save = ops.bn_fwd(xd, gamma.detach(), beta.detach(), rm, rv, True, act=act, residual=_nhwc(res) if res is not None else None,
                         out_pad=pad)
    e = dict(y=_rel_dev(_nchw(y, c), y_ref.detach()), rm=_rel_dev(rm, rm_ref), rv=_rel_dev(rv, rv_ref))
    dyf = ops.reflect_fold(_nhwc(dy), pad)
    dx, dgm, dbt = ops.bn_bwd(xd, dyf, gamma.detach(), beta.detach(), save, True, act=act)
    torch.cuda.synchronize()
    e.update(dx=_rel_dev(_nchw(dx, c), dx_ref), dgamma=_rel_dev(dgm, dg_ref), dbeta=_rel_dev(dbt, db_ref))
    _report(f"c2_bn/c{c}_h{h}", **e)
    assert e["y"] < TOL_BF16 and e["rm"] < 1e-3 and e["rv"] < 1e-3, e
    assert e["dx"] < 2e-2 and e["dgamma"] < 5e-3 and e["dbeta"] < 5e-3, e


def test_c2_reflect_pad_fold_545mb():
    """dei2i_reflect_pad / _fold on the largest tensor of the step (32 x 258 x 258 x 128 BF16 = 545 MB, byte offsets > 2^31)."""
    from de_i2i_gan_b200 import ops
    dev = torch.device("cuda", 0)
    g = torch.Generator(device="cuda").manual_seed(8)
    x = torch.randn(32, 128, 256, 256, generator=g, device=dev).bfloat16().float().requires_grad_(True)
    xp_ref = F.pad(x, (1, 1, 1, 1), mode="reflect")
    xp = ops.reflect_pad(_nhwc(x.detach()), 1)
    assert torch.equal(_nchw(xp, 128), xp_ref.detach())
    d = torch.randn(xp_ref.shape, generator=g, device=dev).bfloat16().float()
    (dx_ref,) = torch.autograd.grad(xp_ref, x, d)
    dx = ops.reflect_fold(_nhwc(d), 1)
    torch.cuda.synchronize()
    assert _rel_dev(_nchw(dx, 128), dx_ref) < TOL_BF16


def test_c2_rowpacked_stem_256_batch32():
    """Generator stem 3->64 7x7 reflect with spectral norm at 256x256, batch 32 through rowpack + conv_tma (7x1 over 64 channels)
    vs nn.Conv2d in FP32 on the GPU: output, input gradient, weight gradient."""
    import copy
    from de_i2i_gan_b200 import functional as Fn, ops
    torch.manual_seed(3)
    conv = torch.nn.Conv2d(3, 64, 7, padding="same", padding_mode="reflect", bias=False)
    with torch.no_grad():
        conv.weight.copy_(conv.weight.bfloat16().float())
    conv = torch.nn.utils.spectral_norm(conv).cuda()
    ours = copy.deepcopy(conv)
    g = torch.Generator(device="cuda").manual_seed(4)
    x = torch.randn(32, 3, 256, 256, generator=g, device="cuda").bfloat16().float().requires_grad_(True)
    y_ref = conv(x)
    dy = torch.randn(y_ref.shape, generator=g, device="cuda").bfloat16().float()
    dx_ref, dw_ref = torch.autograd.grad(y_ref, (x, conv.weight_orig), dy)
    spec = ops.ConvSpec(3, 64, 7, 1, 3, "reflect")
    xd = x.detach().clone().requires_grad_(True)
    with _KernelLog() as kl:
        y = Fn.rowpack_conv(xd, ours, spec, True)
        dxd, dwd = torch.autograd.grad(y.t, (xd, ours.weight_orig), _nhwc(dy))
    torch.cuda.synchronize()
    e = dict(y=_rel_dev(_nchw(y.t, 64), y_ref.detach()), dx=_rel_dev(dxd, dx_ref), dw=_rel_dev(dwd, dw_ref))
    _report("c2_rowpack_stem", **e, kernels=kl.variants())
    assert e["y"] < TOL_BF16 and e["dx"] < TOL_BF16 and e["dw"] < TOL_BF16, e
    assert any(k.startswith("conv_ring") for k in kl.variants()) and any(k.startswith("wgrad_ring") for k in kl.variants())
