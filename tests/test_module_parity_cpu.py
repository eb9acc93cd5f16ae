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
    """What the reference itself cannot run is refused, not emulated: skip connections (architecture.py:451 TODO) and SEAN's
    style distillation (normalization.py:189-190 raises on the model's own 3-D embeddings)."""
    for over in (dict(skip_conn=True), dict(style_norm_block_type="sean", style_distill=True)):
        opt = ref_shim.default_opt(image_size=128, ngf=16, ndf=8, hidden_nc=32, **over)
        with pytest.raises(NotImplementedError):
            _ours(opt)


@pytest.mark.skipif(not ref_shim.available(), reason="/root/reference not present")
@pytest.mark.parametrize("kind,alpha", [("adain", 1), ("adain", 0), ("sean", 1), ("sean", None)])
def test_style_variants_same_seed_same_parameters(kind, alpha):
    """SEAN / AdaIN generators and the StyleExtractor: identical state dicts (keys, order, values) for the same seed."""
    ref_shim.import_reference()
    from models.networks.extractor import StyleExtractor as RefE  # type: ignore
    from models.networks.generator import DefectGanGenerator as RefG  # type: ignore
    from de_i2i_gan_b200.models.networks.extractor import StyleExtractor
    from de_i2i_gan_b200.models.networks.generator import DefectGanGenerator
    opt = ref_shim.default_opt(image_size=64, ngf=8, ndf=4, num_res=2, hidden_nc=16, style_norm_block_type=kind, sean_alpha=alpha,
                               embed_nc=24, latent_dim=16)
    pairs = [(RefG, DefectGanGenerator)] + ([(RefE, StyleExtractor)] if kind == "adain" else [])
    for ref_cls, our_cls in pairs:
        torch.manual_seed(5)
        a = ref_cls(opt)
        a.init_weights("normal", 0.02)
        torch.manual_seed(5)
        b = our_cls(opt)
        b.init_weights("normal", 0.02)
        sa, sb = a.state_dict(), b.state_dict()
        assert list(sa) == list(sb)
        assert all(torch.equal(sa[k], sb[k]) for k in sa)
        b.load_state_dict(sa)
        a.load_state_dict(sb)
