"""DiffAugment: the oracle restatement against the unmodified reference (live import, build container only) and the RNG
contract of the product module (same torch calls, shapes and order as defectGAN/utils/diffaug.py:22-58) - CPU only."""
import importlib.util
import os

import pytest
import torch

from oracle import diffaug_ref as O

REF = "/root/reference/defectGAN/utils/diffaug.py"
POLICIES = ["color", "translation", "cutout", "color,translation,cutout", "cutout,color"]


@pytest.mark.skipif(not os.path.exists(REF), reason="reference tree not present (GPU box)")
@pytest.mark.parametrize("policy", POLICIES)
def test_oracle_matches_reference(policy):
    spec = importlib.util.spec_from_file_location("ref_diffaug", REF)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    for h, w in ((16, 16), (15, 20), (32, 24)):
        x = torch.randn(4, 3, h, w, generator=torch.Generator().manual_seed(h))
        torch.manual_seed(5)
        y_ref = ref.DiffAugment(x, policy)
        torch.manual_seed(5)
        y = O.apply(x, policy, O.draw(x, policy))
        assert torch.allclose(y, y_ref, atol=1e-6), (policy, h, w)


@pytest.mark.parametrize("policy", POLICIES[:5])
def test_product_draws_like_the_oracle(policy):
    """de_i2i_gan_b200.utils.diffaug._draw consumes the torch RNG exactly like the reference functions do."""
    from de_i2i_gan_b200.utils import diffaug as D
    x = torch.randn(3, 3, 20, 28)
    torch.manual_seed(9)
    want = O.draw(x, policy)
    torch.manual_seed(9)
    got = {nm: D._draw(nm, x) for nm in policy.split(",")}
    for nm in want:
        for k in want[nm]:
            assert torch.equal(want[nm][k], got[nm][k]), (nm, k)
    p = D.pack_params(3, x.device, got)
    assert p.shape == (3, 8)
    assert [list(r) for r in D._groups(["cutout", "color", "translation"])] == [["cutout"], ["color", "translation"]]
