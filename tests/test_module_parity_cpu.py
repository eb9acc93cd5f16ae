"""CPU-only drop-in checks of the mirrored nn.Modules: same state-dict keys / shapes / dtypes / order as the reference,
same parameters for the same seed (init RNG order), reference checkpoints load.  Needs /root/reference for the live
comparison; the key layout is also checked against the oracle's reference-format random states (travels)."""
import pytest
import torch

from oracle import ref_shim, defectgan_ref as R


def _ours(opt):
    from de_i2i_gan_b200.models.networks.generator import DefectGanGenerator
    from de_i2i_gan_b200.models.networks.discriminator import DefectGanDiscriminator
    return DefectGanGenerator(opt), DefectGanDiscriminator(opt)


def test_state_dict_layout_matches_reference_format():
    opt = ref_shim.default_opt(image_size=128, ngf=16, ndf=8, hidden_nc=32)
    G, D = _ours(opt)
    rg = R.random_generator_state(ngf=16, hidden_nc=32)
    rd = R.random_discriminator_state(ndf=8, image_size=128)
    gs, ds = G.state_dict(), D.state_dict()
    assert list(gs) == list(rg) and all(gs[k].shape == rg[k].shape and gs[k].dtype == rg[k].dtype for k in rg)
    assert list(ds) == list(rd) and all(ds[k].shape == rd[k].shape for k in rd)
    assert [n for n, _ in G.named_parameters()] == R.trainable_keys(rg)
    G.load_state_dict(rg)  # strict
    D.load_state_dict(rd)


@pytest.mark.skipif(not ref_shim.available(), reason="/root/reference not present")
def test_same_seed_gives_same_parameters_as_reference():
    ref = ref_shim.import_reference()
    opt = ref_shim.default_opt(image_size=128, ngf=16, ndf=8, hidden_nc=32)
    torch.manual_seed(123)
    Gr, Dr = ref.DefectGanGenerator(opt), ref.DefectGanDiscriminator(opt)
    Gr.init_weights("normal", 0.02)
    Dr.init_weights("normal", 0.02)
    torch.manual_seed(123)
    G, D = _ours(opt)
    G.init_weights("normal", 0.02)
    D.init_weights("normal", 0.02)
    for a, b in ((Gr, G), (Dr, D)):
        sa, sb = a.state_dict(), b.state_dict()
        assert list(sa) == list(sb)
        for k in sa:
            assert torch.equal(sa[k], sb[k]), k
    # a reference checkpoint loads into the mirrored module and vice versa
    G.load_state_dict(Gr.state_dict())
    Gr.load_state_dict(G.state_dict())


def test_unsupported_variants_fail_loudly():
    opt = ref_shim.default_opt(image_size=128, ngf=16, ndf=8, hidden_nc=32, style_norm_block_type="sean")
    with pytest.raises(NotImplementedError):
        _ours(opt)
