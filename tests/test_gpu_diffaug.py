"""DiffAugment on the fused kernels (csrc/augment.cu) vs the oracle restatement (oracle/diffaug_ref.py, pinned to the reference):
forward and backward with the same drawn parameters, FP32 - north-star tolerance 1e-3 relative (measured ~1e-6)."""
import pytest
import torch

from conftest import rel_err

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("policy", ["color", "translation", "cutout", "color,translation,cutout", "cutout,color", "translation,cutout"])
@pytest.mark.parametrize("n,h,w", [(4, 32, 32), (3, 15, 20), (32, 256, 256)])
def test_diffaug_fwd_bwd(policy, n, h, w):
    from oracle import diffaug_ref as O
    from de_i2i_gan_b200.utils.diffaug import DiffAugment
    g = torch.Generator().manual_seed(3)
    x = torch.randn(n, 3, h, w, generator=g)
    torch.manual_seed(11)
    drawn = O.draw(x, policy)
    xr = x.clone().requires_grad_(True)
    y_ref = O.apply(xr, policy, drawn)
    dy = torch.randn(y_ref.shape, generator=g)
    (dx_ref,) = torch.autograd.grad(y_ref, xr, dy)
    xd = x.cuda().requires_grad_(True)
    y = DiffAugment(xd, policy, _params={k: {kk: vv.cuda() for kk, vv in v.items()} for k, v in drawn.items()})
    (dx,) = torch.autograd.grad(y, xd, dy.cuda())
    torch.cuda.synchronize()
    assert rel_err(y, y_ref) < 1e-5 and rel_err(dx, dx_ref) < 1e-5, (rel_err(y, y_ref), rel_err(dx, dx_ref))


def test_diffaug_rng_contract_on_device():
    """Without injected parameters the module draws on the device with the reference's calls: same seed -> same augmentation twice,
    and (policy='') is the identity."""
    from de_i2i_gan_b200.utils.diffaug import DiffAugment
    x = torch.randn(4, 3, 32, 32, device="cuda")
    assert DiffAugment(x, "") is x
    torch.manual_seed(3)
    a = DiffAugment(x, "color,translation,cutout")
    torch.manual_seed(3)
    b = DiffAugment(x, "color,translation,cutout")
    assert torch.equal(a, b) and not torch.equal(a, x)


def test_defectgan_model_with_diff_aug_steps():
    """DefectGanModel with --diff_aug color,translation,cutout (train_mtvec.py:95): both loss graphs run, gradients reach G through
    the augmentation and D."""
    from de_i2i_gan_b200 import options
    from de_i2i_gan_b200.models.defectgan_model import DefectGanModel
    from de_i2i_gan_b200.synthetic import synthetic_batch
    opt = options.readme_recipe(image_size=128, batch_size=2, ngf=16, ndf=8, hidden_nc=32, diff_aug="color,translation,cutout")
    opt.device = torch.device("cuda", 0)
    torch.manual_seed(1)
    model = DefectGanModel(opt)
    model.init_weights()
    bg, labels, df = [t.cuda() for t in synthetic_batch(2, 128, seed=0)]
    gan, clf = model("discriminator", bg, labels, df)
    (gan + clf).backward()
    assert all(torch.isfinite(p.grad).all() for p in model.netD.parameters() if p.grad is not None)
    losses = model("generator", bg, labels, df)
    sum(losses).backward()
    g = [p.grad for p in model.netG.parameters() if p.grad is not None]
    assert g and all(torch.isfinite(t).all() for t in g) and any(t.abs().max() > 0 for t in g)
