"""MAE pre-training path on the CUDA kernels vs fixtures from the UNMODIFIED reference (oracle/make_golden_mae.py):
DefectGanModel('mae_discriminator' / 'mae_generator') for every mask-token type - same state dicts, images, labels, the mask that
the same host seed draws, injected noise.  BF16 tolerances as tests/test_gpu_golden.py."""
import os

import pytest
import torch

from conftest import GOLDEN, rel_err
from test_gpu_golden import _clear_noise, _grad_check, _opt, _queue_noise

pytestmark = pytest.mark.gpu
TOL = 2e-2


def _model(t, case, first):
    from de_i2i_gan_b200.models.defectgan_model import DefectGanModel
    opt = _opt(case["opt"])
    opt.device = torch.device("cuda", 0)
    model = DefectGanModel(opt)
    model.netG.load_state_dict(case["g_state"] or first["g_state"])
    model.netD.load_state_dict(case["d_state"] or first["d_state"])
    model.netG.cuda()
    model.netD.cuda()
    if case["token"] is not None:
        with torch.no_grad():
            model.mask_token.mask_token.copy_(case["token"])
    return model


@pytest.mark.parametrize("ttype", ["position", "zero", "mean", "vector", "scalar", "full"])
def test_mae_losses_and_gradients(ttype):
    t = torch.load(os.path.join(GOLDEN, "mae_small.pt"), weights_only=False)
    case, first = t["cases"][ttype], t["cases"]["position"]
    model = _model(t, case, first)
    imgs, labels = t["imgs"].cuda(), t["labels"].cuda()
    try:
        q = _queue_noise(case["d"]["noises"])
        torch.manual_seed(77)  # the host RNG draws the same shifted mask as in the fixture run (utils/util.py:60-71)
        gan_d, clf_d = model("mae_discriminator", imgs, labels)
        (gan_d + 3 * clf_d).backward()
        assert not q
        if case["d"]["grads"] is not None:
            _grad_check(model.netD.named_parameters(), case["d"]["grads"], f"mae D {ttype}")
        for p in model.netD.parameters():
            p.grad = None
            p.requires_grad_(False)
        q = _queue_noise(case["g"]["noises"])
        torch.manual_seed(78)
        rec, gan, clf = model("mae_generator", imgs, labels)
        (gan + 10 * rec + clf).backward()
        assert not q
    finally:
        _clear_noise()
    got = dict(gan_d=gan_d.item(), clf_d=clf_d.item(), rec=rec.item(), gan=gan.item(), clf=clf.item())
    want = dict(gan_d=case["d"]["gan"], clf_d=case["d"]["clf"], rec=case["g"]["rec"], gan=case["g"]["gan"], clf=case["g"]["clf"])
    for k in want:
        assert abs(got[k] - want[k]) / max(abs(want[k]), 1e-3) < TOL, (ttype, k, got[k], want[k])
    if case["g"]["grads"] is not None:
        _grad_check(model.netG.named_parameters(), case["g"]["grads"], f"mae G {ttype}")
    if case["g"]["token_grad"] is not None:
        tg, ref = model.mask_token.mask_token.grad, case["g"]["token_grad"]
        assert tg is not None
        # the token gradient is the INPUT gradient of the whole generator - the end of the ~30-layer BF16 backward chain (the blend
        # kernel itself is exact: test_mask_blend_matches_reference_formula); measured 0.03 - 0.11 relative L2, cosine > 0.99
        a, b = tg.detach().float().cpu().flatten(), ref.flatten()
        rel, cos = ((a - b).norm() / b.norm()).item(), (torch.dot(a, b) / (a.norm() * b.norm())).item()
        assert rel < 0.15 and cos > 0.985, (ttype, rel, cos)


def test_mask_blend_matches_reference_formula():
    """dei2i_mask_token_fwd / _bwd vs MaskToken.forward's own formula on the materialised mask (FP32, exact to rounding)."""
    from types import SimpleNamespace
    from de_i2i_gan_b200.models.networks.architecture import MaskToken
    from de_i2i_gan_b200.utils.util import draw_shifted_mask, expand_mask
    g = torch.Generator().manual_seed(2)
    imgs = torch.randn(3, 3, 64, 64, generator=g)
    for ttype in ("zero", "mean", "scalar", "vector", "position", "full"):
        opt = SimpleNamespace(mask_token_type=ttype, mask_ratio=0.75, input_nc=3, image_size=64)
        mt = MaskToken(opt).cuda()
        if isinstance(mt.mask_token, torch.nn.Parameter):
            with torch.no_grad():
                mt.mask_token.copy_(torch.randn(mt.mask_token.shape, generator=g))
        torch.manual_seed(4)
        coarse, hs, ws = draw_shifted_mask(imgs.size(), 8, 0.75)
        mask = expand_mask(coarse, hs, ws, imgs.size(), 8)
        out, m = mt.blend(imgs.cuda(), coarse, hs, ws, 8)
        ref_mod = MaskToken(opt)
        if isinstance(ref_mod.mask_token, torch.nn.Parameter):
            ref_mod.mask_token.data.copy_(mt.mask_token.detach().cpu())
        ref = ref_mod(imgs, mask)
        assert torch.equal(m.cpu(), mask) and rel_err(out, ref) < 1e-6, ttype
        if isinstance(mt.mask_token, torch.nn.Parameter):
            dy = torch.randn(out.shape, generator=g)
            (gt,) = torch.autograd.grad(out, mt.mask_token, dy.cuda())
            (gr,) = torch.autograd.grad(ref, ref_mod.mask_token, dy)
            assert rel_err(gt, gr) < 1e-5, (ttype, rel_err(gt, gr))


def test_mae_trainer_steps_with_adamw():
    """MAETrainer (mae_trainer.py:121-158) with the pre-training defaults (AdamW, 'position' token): two iterations run, losses are
    finite, generator, discriminator and mask token all move."""
    from pathlib import Path
    from de_i2i_gan_b200 import options
    from de_i2i_gan_b200.synthetic import synthetic_batch
    from de_i2i_gan_b200.trainers.mae_trainer import MAETrainer
    torch.manual_seed(5)
    opt = options.pretrain_options(image_size=128, batch_size=2, ngf=16, ndf=8, hidden_nc=32, add_noise=True, use_spectral=True,
                                   num_epochs=2, iters_per_epoch=4, ckpt_dir=Path("/tmp/dei2i_mae_test"), name="t")
    opt.device = torch.device("cuda", 0)
    tr = MAETrainer(opt, data_types=("fusion",))
    imgs, labels, _ = synthetic_batch(2, 128, seed=0)
    g0 = [p.detach().clone() for p in tr.model.netG.parameters()]
    d0 = [p.detach().clone() for p in tr.model.netD.parameters()]
    for _ in range(2):
        tr._train_discriminator_once(imgs, labels)
        tr._train_generator_once(imgs, labels)
    vals = [v for t in tr.losses.values() for lst in t.values() for v in lst]
    assert len(vals) == 10 and all(v == v and abs(v) < 1e3 for v in vals), tr.losses
    assert any(not torch.equal(a, b) for a, b in zip(g0, tr.model.netG.parameters()))
    assert any(not torch.equal(a, b) for a, b in zip(d0, tr.model.netD.parameters()))
    assert tr.model.mask_token.mask_token.abs().max().item() > 0  # the token is initialised to zeros and is in G's optimizer
