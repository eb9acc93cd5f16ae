"""SEAN / AdaIN style variants, host side, no GPU: the mirrored module trees carry exactly the reference's
state-dict keys and shapes (tests/golden/style_small.pt, written by the unmodified reference via oracle/make_golden_style.py), the
embedding draw follows the reference's host RNG calls, and the modes the reference itself cannot run are refused loudly."""
import os
import random
from types import SimpleNamespace

import pytest
import torch

from conftest import GOLDEN


def _load():
    return torch.load(os.path.join(GOLDEN, "style_small.pt"), weights_only=False)


def _model(case, fixture, tmp_path):
    from de_i2i_gan_b200.models.defectgan_model import DefectGanModel
    opt = SimpleNamespace(**case["opt"])
    opt.device = torch.device("cpu")
    opt.embed_path = None
    if opt.style_norm_block_type == "sean" and opt.sean_alpha != 0:
        opt.embed_path = tmp_path / "emb.pt"
        torch.save(fixture["embeddings"], opt.embed_path)
    return DefectGanModel(opt), opt


@pytest.mark.parametrize("name", ["adain_1", "adain_0", "sean_1", "sean_0"])
def test_state_dict_keys_and_shapes_match_reference(name, tmp_path):
    fx = _load()
    case = fx["cases"][name]
    model, _ = _model(case, fx, tmp_path)
    assert sorted(model.networks) == sorted(case["keys"])
    for net_name, keys in case["keys"].items():
        st = model.networks[net_name].state_dict()
        assert list(st) == keys, (net_name, [k for k in keys if k not in st][:5], [k for k in st if k not in keys][:5])
        ref = case["states"][net_name]
        if isinstance(ref, str):
            ref = fx["cases"][ref]["states"][net_name]
        for k, v in ref.items():
            assert tuple(st[k].shape) == tuple(v.shape), (net_name, k)
        missing, unexpected = model.networks[net_name].load_state_dict(ref, strict=False)
        assert not unexpected and all(m.rsplit(".", 1)[-1].startswith(("mean_", "std_")) for m in missing), (missing[:5], unexpected[:5])


def test_style_embeds_follow_reference_draws(tmp_path):
    """defectgan_model.py:394-411: `random.choices(pool, k=num_embeds)` per sample, zeros when the label combination has no pool."""
    fx = _load()
    model, opt = _model(fx["cases"]["sean_1"], fx, tmp_path)
    labels = fx["labels"]
    random.seed(3)
    got = model._get_style_embeds(labels)
    random.seed(3)
    want = []
    for row in labels.int().tolist():
        pool = fx["embeddings"][tuple(row)]
        want.append(torch.stack(random.choices(pool, k=opt.num_embeds)) if pool else torch.zeros(opt.num_embeds, opt.embed_nc))
    assert got.shape == (labels.size(0), opt.num_embeds, opt.embed_nc)
    assert torch.equal(got, torch.stack(want))
    assert not got[1].any() and got[0].any()  # the fixture's second label row has no stored embeddings
    model.opt.sean_alpha = 0
    assert model._get_style_embeds(labels) is None


def test_sean_running_stats_bookkeeping():
    """normalization.py:106-125 incl. the reference's crosswise store (mean <- std, std <- mean), which checkpoints depend on."""
    from de_i2i_gan_b200.models.networks.normalization import SEAN, label_to_str
    from torch import nn
    s = SEAN(8, 4, 3, hidden_nc=5, norm_layer=nn.InstanceNorm2d)
    assert len([k for k in s.state_dict() if k.startswith("mean_")]) == 8 and label_to_str((1, 0, 1)) == "0-2"
    feats = [torch.randn(5) for _ in range(6)]
    s.embeds[(1, 0, 1)] = list(feats)
    s.update_stats()
    f = torch.stack(feats)
    assert torch.allclose(getattr(s, "mean_0-2"), (f.var(dim=0) + 1e-5).sqrt())
    assert torch.allclose(getattr(s, "std_0-2"), f.mean(dim=0))
    assert not getattr(s, "mean_1").any()


def test_unrunnable_reference_modes_are_refused():
    from de_i2i_gan_b200.models.networks.normalization import SEAN, AdaIN
    from torch import nn
    with pytest.raises(NotImplementedError):
        SEAN(8, 4, 3, hidden_nc=5, norm_layer=nn.InstanceNorm2d, style_distill=True)
    with pytest.raises(NotImplementedError):
        AdaIN(4, hidden_nc=5, norm_layer=nn.BatchNorm2d)
