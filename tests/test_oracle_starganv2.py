"""oracle/starganv2_ref.py (CPU restatement of the StarGAN-v2 path, adain / w_hpf = 0) against tests/golden/starganv2_small.pt,
produced from the UNMODIFIED reference (stargan-v2/core/{model,solver}.py) by oracle/make_golden_starganv2.py."""
import os
from types import SimpleNamespace

import torch

from oracle import defectgan_ref as R0
from oracle import starganv2_ref as R
from conftest import rel_err, GOLDEN

TOL = 1e-5


def _load():
    t = torch.load(os.path.join(GOLDEN, "starganv2_small.pt"), weights_only=False)
    c = t["cfg"]
    f = R.mapping_state(t["mapping_seed"], c["latent_dim"], c["style_dim"], c["num_domains"])
    assert abs(float(sum(v.double().sum() for v in f.values())) - t["mapping_checksum"]) < 1e-6 * max(1.0, abs(t["mapping_checksum"]))
    nets = SimpleNamespace(generator=R0.clone_state(t["states"]["generator"]), mapping_network=f,
                           style_encoder=R0.clone_state(t["states"]["style_encoder"]),
                           discriminator=R0.clone_state(t["states"]["discriminator"]))
    args = SimpleNamespace(lambda_reg=c["lambda_reg"], lambda_sty=c["lambda_sty"], lambda_ds=c["lambda_ds"], lambda_cyc=c["lambda_cyc"])
    return t, nets, args


def _req(nets):
    for st in vars(nets).values():
        for v in st.values():
            if v.is_floating_point():
                v.requires_grad_(True)


def test_networks_forward():
    t, nets, _ = _load()
    with torch.no_grad():
        s_map = R.mapping_forward(nets.mapping_network, t["z_trg"], t["y_trg"])
        assert rel_err(s_map, t["fwd"]["s_map"]) < TOL
        assert rel_err(R.style_encoder_forward(nets.style_encoder, t["x_ref"], t["y_trg"]), t["fwd"]["s_enc"]) < TOL
        assert rel_err(R.generator_forward(nets.generator, t["x_real"], s_map), t["fwd"]["x_fake"]) < TOL
        assert rel_err(R.discriminator_forward(nets.discriminator, t["x_real"], t["y_org"]), t["fwd"]["d_out"]) < TOL


def test_d_loss_with_r1_double_backward():
    t, nets, args = _load()
    _req(nets)
    loss, log = R.compute_d_loss(nets, args, t["x_real"], t["y_org"], t["y_trg"], z_trg=t["z_trg"])
    ref = t["res"]["d_latent"]
    assert abs(loss.item() - ref["loss"]) < 1e-5 and abs(log.reg - ref["log"]["reg"]) < 1e-5 * max(1.0, ref["log"]["reg"])
    loss.backward()
    for n, g in ref["grads"].items():
        assert rel_err(nets.discriminator[n].grad, g) < 1e-3, n
    loss, _ = R.compute_d_loss(nets, args, t["x_real"], t["y_org"], t["y_trg"], x_ref=t["x_ref"])
    assert abs(loss.item() - t["res"]["d_ref"]["loss"]) < 1e-5


def test_g_loss():
    t, nets, args = _load()
    _req(nets)
    loss, log = R.compute_g_loss(nets, args, t["x_real"], t["y_org"], t["y_trg"], z_trgs=[t["z_trg"], t["z_trg2"]])
    ref = t["res"]["g_latent"]
    assert abs(loss.item() - ref["loss"]) < 1e-4, (loss.item(), ref["loss"])
    for k in ("adv", "sty", "ds", "cyc"):
        assert abs(getattr(log, k) - ref["log"][k]) < 1e-4, k
    loss.backward()
    for st, key in ((nets.generator, "grads_g"), (nets.style_encoder, "grads_e"), (nets.mapping_network, "grads_f")):
        gmax = max(g.abs().max().item() for g in ref[key].values())
        for n, g in ref[key].items():  # biases in front of an instance norm have exactly-zero gradients (rounding noise only)
            assert (st[n].grad - g).abs().max().item() < 2e-3 * g.abs().max().item() + 1e-4 * gmax, (key, n)
    loss, _ = R.compute_g_loss(nets, args, t["x_real"], t["y_org"], t["y_trg"], x_refs=[t["x_ref"], t["x_ref2"]])
    assert abs(loss.item() - t["res"]["g_ref"]["loss"]) < 1e-4
