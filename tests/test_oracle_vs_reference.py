"""Live pin of the oracle restatement against the unmodified reference (build container only: skipped where
/root/reference is absent, e.g. on the GPU box)."""
import pytest
import torch

from oracle import ref_shim, defectgan_ref as R
from conftest import rel_err

pytestmark = pytest.mark.skipif(not ref_shim.available(), reason="/root/reference not present")


def _nets(image_size):
    ref = ref_shim.import_reference()
    torch.manual_seed(123)
    opt = ref_shim.default_opt(image_size=image_size, ngf=16, ndf=8, hidden_nc=32)
    G, D = ref.DefectGanGenerator(opt), ref.DefectGanDiscriminator(opt)
    G.init_weights("normal", 0.02)
    D.init_weights("normal", 0.02)
    for n, p in G.named_parameters():
        if "noise" in n:
            p.data.fill_(0.1)
    return G, D


def test_state_layout_matches_reference_state_dict():
    G, D = _nets(128)
    rg = R.random_generator_state(ngf=16, hidden_nc=32)
    rd = R.random_discriminator_state(ndf=8, image_size=128)
    gs, ds = G.state_dict(), D.state_dict()
    assert list(rg) == list(gs) and all(rg[k].shape == gs[k].shape and rg[k].dtype == gs[k].dtype for k in gs)
    assert list(rd) == list(ds) and all(rd[k].shape == ds[k].shape for k in ds)
    assert R.trainable_keys(gs) == [n for n, _ in G.named_parameters()]


def test_generator_live():
    G, _ = _nets(128)
    st = R.clone_state(G.state_dict())
    for k in R.trainable_keys(st):
        st[k].requires_grad_(True)
    bg, lab, _ = R.synthetic_batch(2, 64, seed=3)
    G.train()
    torch.manual_seed(7)
    o_ref, p_ref = G(bg, lab.reshape(2, 6, 1, 1))
    torch.manual_seed(7)  # the restatement draws noise from the same RNG in the same order
    o, p = R.generator_forward(st, bg, lab, True)
    assert rel_err(o, o_ref) < 1e-5 and rel_err(p, p_ref) < 1e-5
    (o_ref.square().mean() + p_ref.mean()).backward()
    (o.square().mean() + p.mean()).backward()
    gmax = max(q.grad.abs().max().item() for q in G.parameters() if q.grad is not None)
    for n, q in G.named_parameters():
        if q.grad is None:
            assert st[n].grad is None, n
        else:
            assert (q.grad - st[n].grad).abs().max().item() < 1e-4 * q.grad.abs().max().item() + 1e-7 * gmax, n
    sd = G.state_dict()
    for k in sd:
        if k.endswith(("_u", "_v", "running_mean", "running_var")):
            assert rel_err(st[k], sd[k]) < 1e-5, k


def test_discriminator_live():
    _, D = _nets(128)
    st = R.clone_state(D.state_dict())
    for k in R.trainable_keys(st):
        st[k].requires_grad_(True)
    x, _, _ = R.synthetic_batch(2, 128, seed=4)
    for training in (True, False):
        D.train(training)
        s_ref, c_ref = D(x)
        s, c = R.discriminator_forward(st, x, training)
        assert rel_err(s, s_ref) < 1e-5 and rel_err(c, c_ref) < 1e-5
