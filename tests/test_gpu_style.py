"""SEAN / AdaIN style variants of DefectGAN on the CUDA path against the UNMODIFIED reference on CPU in FP32
(tests/golden/style_small.pt, oracle/make_golden_style.py): one D-loss and one G-loss through DefectGanModel per variant, same state
dicts, inputs, injected NoiseInjection maps and host RNG protocol (`random.seed(s); torch.manual_seed(s)` before each call: the
StyleExtractor's host noise and the embedding choices are the first host draws of a call on both sides).

Tolerances as tests/test_gpu_golden.py: losses and outputs 2e-2 relative, whole-network gradients global rel-L2 < 8e-2 and cosine
> 0.9 on every parameter carrying > 1e-3 of the gradient norm."""
import os
import random
from types import SimpleNamespace

import pytest
import torch

from conftest import GOLDEN, rel_err
from test_gpu_golden import TOL_ACT, _clear_noise, _grad_check, _queue_noise, _report

pytestmark = pytest.mark.gpu
W = dict(clf_d=2, clf_g=5, rec=5, sd_cyc=5, sd_con=1)


def _build(name, tmp_path):
    from de_i2i_gan_b200.models.defectgan_model import DefectGanModel
    fx = torch.load(os.path.join(GOLDEN, "style_small.pt"), weights_only=False)
    case = fx["cases"][name]
    opt = SimpleNamespace(**case["opt"])
    opt.device = torch.device("cuda", 0)
    opt.embed_path = None
    if opt.style_norm_block_type == "sean" and opt.sean_alpha != 0:
        opt.embed_path = tmp_path / "emb.pt"
        torch.save(fx["embeddings"], opt.embed_path)
    model = DefectGanModel(opt)
    for net_name, net in model.networks.items():
        st = case["states"][net_name]
        if isinstance(st, str):
            st = fx["cases"][st]["states"][net_name]
        net.load_state_dict(st, strict=False)  # SEAN's all-zero running-statistics buffers are not stored (key parity: test_style_cpu.py)
        net.cuda()
    return fx, case, model


def _seed(s):
    random.seed(s)
    torch.manual_seed(s)


@pytest.mark.parametrize("name", ["adain_1", "adain_0", "sean_1", "sean_0"])
def test_style_variant_losses_and_gradients(name, tmp_path):
    fx, case, model = _build(name, tmp_path)
    bg, df, lab = fx["bg"].cuda(), fx["df"].cuda(), fx["labels"].cuda()
    ref = case["losses"]
    try:
        q = _queue_noise(case["noises_d"])
        _seed(7)
        gan_d, clf_d = model("discriminator", bg, lab, df)
        model.netD.zero_grad()
        (gan_d + clf_d * W["clf_d"]).backward()
        assert not q, "noise draws out of order / count"
        # the discriminator is not what varies here (its own parity: test_gpu_golden.py); the fixture keeps its first and last
        # parameter only, i.e. the two ends of the BF16 backward chain: direction check
        d_worst = 1.0
        for n, p in model.netD.named_parameters():
            if n in case["d_grads"]:
                a, b = p.grad.detach().float().cpu().flatten(), case["d_grads"][n].flatten()
                d_worst = min(d_worst, (torch.dot(a, b) / (a.norm() * b.norm() + 1e-30)).item())
        assert d_worst > 0.9, (name, "D gradient cosine", d_worst)
        q = _queue_noise(case["noises_g"])
        for net in model.networks.values():
            net.zero_grad()
        _seed(8)
        gan, clf, rec, cyc, con = model("generator", bg, lab, df)
        (gan + clf * W["clf_g"] + rec * W["rec"] + cyc * W["sd_cyc"] + con * W["sd_con"]).backward()
        assert not q
        got = dict(gan_d=gan_d.item(), clf_d=clf_d.item(), gan_g=gan.item(), clf_g=clf.item(), rec=rec.item(), cyc=cyc.item(), con=con.item())
        worst_loss = 0.0
        for k, v in ref.items():
            err = abs(got[k] - v) / max(abs(v), 1e-3)
            worst_loss = max(worst_loss, err)
            assert err < TOL_ACT, (name, k, got[k], v)
        g_worst = _grad_check(model.netG.named_parameters(), case["g_grads"], name + "/G")
        e_worst = None
        if case["e_grads"]:
            # the extractor sits behind D, two chained generator passes and its own ten convolutions + InstanceNorms: the longest BF16
            # backward chain of the repository (measured 0.097 at these 4..32-channel widths) - 0.15, direction per parameter as ever
            e_worst = _grad_check([(n, p) for n, p in model.netE.named_parameters() if n in case["e_grads"]], case["e_grads"], name + "/E",
                                  l2_bar=0.15)
        q = _queue_noise(case["noises_eval"])
        _seed(9)
        model.netG.eval()
        with torch.no_grad():
            out, prob = model._generate_fake(bg, lab)
        assert not q
    finally:
        _clear_noise()
    e_out, e_prob = rel_err(out, case["out_eval"]), rel_err(prob, case["prob_eval"])
    _report("golden/style_" + name, worst_loss_rel=worst_loss, g_grads=g_worst, d_grads=d_worst, e_grads=e_worst, out_eval=e_out, prob_eval=e_prob)
    assert e_out < TOL_ACT and e_prob < TOL_ACT, (e_out, e_prob)


def test_adain_trainer_steps_extractor(tmp_path):
    """DefectGanTrainer with the AdaIN variant: netE gets its own optimizer and is stepped with the generator
    (defectgan_trainer.py:140-141,162-163); the whole G-step incl. the extractor is captured as one CUDA graph."""
    from pathlib import Path
    from de_i2i_gan_b200.trainers.defectgan_trainer import DefectGanTrainer
    from de_i2i_gan_b200.options import train_options
    fx = torch.load(os.path.join(GOLDEN, "style_small.pt"), weights_only=False)
    net = {k: v for k, v in fx["cases"]["adain_1"]["opt"].items() if k not in ("num_epochs",)}
    opt = train_options(**{k: v for k, v in net.items() if k in vars(train_options())})
    opt.__dict__.update(ckpt_dir=Path(tmp_path), name="t", iters_per_epoch=4, num_epochs=8, num_iters=32, num_critics=1)
    tr = DefectGanTrainer(opt)
    assert sorted(tr.optimizers) == ["D", "E", "G"]
    e0 = [p.detach().clone() for p in tr.model.netE.parameters()]
    g0 = [p.detach().clone() for p in tr.model.netG.parameters()]
    bg, df, lab = fx["bg"], fx["df"], fx["labels"]
    for _ in range(4):  # eager, capture, two replays
        tr._train_discriminator_once(bg, lab, df)
        tr._train_generator_once(bg, lab, df)
    assert all(len(v) == 4 for v in tr.losses["aux"].values())
    assert all(torch.isfinite(torch.tensor(v)).all() for t in tr.losses.values() for v in t.values())
    moved_e = sum((p.detach() - q).abs().sum().item() for p, q in zip(tr.model.netE.parameters(), e0))
    moved_g = sum((p.detach() - q).abs().sum().item() for p, q in zip(tr.model.netG.parameters(), g0))
    assert moved_e > 0 and moved_g > 0
    assert tr.losses["aux"]["rec"][-1] < tr.losses["aux"]["rec"][0]  # four Adam steps on one batch: the reconstruction term falls
