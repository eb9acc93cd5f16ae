"""CPU checks of the MAE host pieces: mask helpers draw exactly like the reference (utils/util.py:48-71), MaskToken keeps the
reference's state-dict key and formula, the fixture generator script is committed with its fixture."""
import os

import pytest
import torch

from conftest import GOLDEN

REF = "/root/reference/defectGAN"


@pytest.mark.skipif(not os.path.isdir(REF), reason="reference tree not present (GPU box)")
def test_mask_helpers_match_reference():
    import importlib.util
    from oracle import ref_shim
    ref_shim.install_stubs()
    spec = importlib.util.spec_from_file_location("ref_util", os.path.join(REF, "utils", "util.py"))
    RU = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(RU)
    from de_i2i_gan_b200.utils import util as U
    for b, h, w, ps, r in ((4, 64, 64, 8, 0.75), (2, 128, 128, 16, 0.5), (3, 32, 64, 8, 0.75)):
        torch.manual_seed(3)
        a = RU.generate_shifted_mask((b, 3, h, w), ps, r)
        torch.manual_seed(3)
        m = U.generate_shifted_mask((b, 3, h, w), ps, r)
        assert torch.equal(a, m)
        torch.manual_seed(4)
        a = RU.generate_mask((b, 3, h, w), ps, r)
        torch.manual_seed(4)
        assert torch.equal(a, U.generate_mask((b, 3, h, w), ps, r))


def test_fixture_masks_reproduce_from_seed():
    """The masks stored in the reference-generated fixture come back from the product's draw with the same host seed."""
    from de_i2i_gan_b200.utils.util import generate_shifted_mask
    t = torch.load(os.path.join(GOLDEN, "mae_small.pt"), weights_only=False)
    case = t["cases"]["position"]
    torch.manual_seed(77)
    assert torch.equal(generate_shifted_mask(t["imgs"].size(), 8, 0.75).bool(), case["d"]["mask"])
    torch.manual_seed(78)
    assert torch.equal(generate_shifted_mask(t["imgs"].size(), 8, 0.75).bool(), case["g"]["mask"])


def test_mask_token_state_dict_and_formula():
    from types import SimpleNamespace
    from de_i2i_gan_b200.models.networks.architecture import MaskToken
    opt = SimpleNamespace(mask_token_type="position", mask_ratio=0.75, input_nc=3, image_size=16)
    mt = MaskToken(opt)
    assert list(mt.state_dict()) == ["mask_token"] and tuple(mt.mask_token.shape) == (1, 1, 16, 16)
    with torch.no_grad():
        mt.mask_token.fill_(0.5)
    x, m = torch.ones(2, 3, 16, 16), torch.zeros(2, 1, 16, 16)
    m[:, :, :8] = 1
    y = mt(x, m)
    assert torch.equal(y[:, :, :8], x[:, :, :8]) and torch.all(y[:, :, 8:] == 0.5)
    mean = MaskToken(SimpleNamespace(mask_token_type="mean", mask_ratio=0.5, input_nc=3, image_size=16))
    y = mean(x, m)
    assert torch.allclose(y[:, :, 8:], torch.ones(2, 3, 8, 16))  # mean(x * m) / ratio = 0.5 / 0.5
