"""oracle/wgan_ref.py (CPU restatement of the WGAN path) against tests/golden/wgan_small.pt, produced from the UNMODIFIED
reference networks by oracle/make_golden_wgan.py - and, where /root/reference exists, against the live reference modules."""
import os

import pytest
import torch

from oracle import defectgan_ref as R
from oracle import ref_shim
from oracle import wgan_ref as W
from conftest import rel_err, GOLDEN

TOL = 1e-5


def _load():
    return torch.load(os.path.join(GOLDEN, "wgan_small.pt"), weights_only=False)


def test_wgan_networks_match_reference_fixture():
    t = _load()
    g, d = R.clone_state(t["g_state"]), R.clone_state(t["d_state"])
    for st in (g, d):
        for k in R.trainable_keys(st):
            st[k].requires_grad_(True)
    fake = W.generator_forward(g, t["z"], True)
    assert rel_err(fake, t["fwd"]["fake"]) < TOL
    fake.square().mean().backward()
    for n, ref in t["fwd"]["g_grads"].items():
        assert rel_err(g[n].grad, ref) < 1e-4, n
    x = t["real"].clone().requires_grad_(True)
    logits = W.discriminator_forward(d, x, True)
    assert rel_err(logits, t["fwd"]["logits"]) < TOL
    logits.sum().backward()
    assert rel_err(x.grad, t["fwd"]["dx"]) < 1e-4
    for n, ref in t["fwd"]["d_grads"].items():
        assert rel_err(d[n].grad, ref) < 1e-4, n
    with torch.no_grad():  # eval mode: running statistics after one training-mode forward each
        assert rel_err(W.generator_forward(g, t["z"], False), t["fwd"]["fake_eval"]) < TOL
        assert rel_err(W.discriminator_forward(d, t["real"], False), t["fwd"]["logits_eval"]) < TOL


def test_wgan_two_steps_match_reference_fixture():
    t = _load()
    tr = W.Trainer(R.clone_state(t["g_state"]), R.clone_state(t["d_state"]), lr=t["cfg"]["lr"], clipping_limit=t["cfg"]["clipping_limit"])
    for it in range(2):
        d_loss = tr.train_discriminator_once(t["real"], t["zs"][2 * it])
        g_loss = tr.train_generator_once(t["zs"][2 * it + 1])
        ref = t["steps"][it]
        assert abs(d_loss - ref["d_loss"]) < 1e-6 + 1e-3 * abs(ref["d_loss"]), (it, d_loss, ref)
        assert abs(g_loss - ref["g_loss"]) < 1e-6 + 1e-3 * abs(ref["g_loss"]), (it, g_loss, ref)
    for st, after in ((tr.g, t["g_state_after"]), (tr.d, t["d_state_after"])):
        for k, ref in after.items():
            if k.endswith("num_batches_tracked"):
                assert int(st[k]) == int(ref), k
            else:
                assert rel_err(st[k], ref) < 1e-2, k  # RMSprop: g / (sqrt(v) + 1e-8) with |g| ~ 1e-6 here amplifies FP32 rounding


@pytest.mark.skipif(not ref_shim.available(), reason="/root/reference is only present in the build container")
def test_wgan_oracle_vs_live_reference():
    ref = ref_shim.import_reference()
    torch.manual_seed(11)
    G = ref.WGanGenerator(num_layers=1, ngf=4, noise_dim=12)
    D = ref.WGanDiscriminator(num_layers=1, ndf=4)
    G.init_weights("normal", 0.02)
    D.init_weights("normal", 0.02)
    g = R.clone_state(G.state_dict())
    d = R.clone_state(D.state_dict())
    z = torch.rand(3, 12, 1, 1)
    x = torch.randn(3, 3, 16, 16)
    G.train(); D.train()
    assert rel_err(W.generator_forward(g, z, True), G(z)) < TOL
    assert rel_err(W.discriminator_forward(d, x, True), D(x)) < TOL
    assert W.num_layers_for(16) == 1
