"""TEST INFRASTRUCTURE ONLY - generates tests/golden/starganv2_small.pt by running the UNMODIFIED reference
(/root/reference/stargan-v2/core/{model,solver}.py through oracle/ref_shim.import_stargan) on CPU in FP32.

    python -m oracle.make_golden_starganv2     # from the repo root, in the build container

Reduced config: img_size 256 (dim_in = 2^14 / 256 = 64 as in configs[3]) with max_conv_dim = 16 so the file stays small,
style_dim 16, latent_dim 8, 3 domains, batch 2, norm_type 'adain', w_hpf 0.  Stored: the four state dicts, the inputs,
network outputs, compute_d_loss / compute_g_loss values (latent and reference branches) and the parameter gradients of the
D loss (incl. the R1 double backward) and of the G loss.
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_shim  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
CFG = dict(img_size=256, style_dim=16, latent_dim=8, num_domains=3, max_conv_dim=16, batch=2, lambda_reg=1.0, lambda_sty=1.0,
           lambda_ds=2.0, lambda_cyc=1.0)


def main():
    sg = ref_shim.import_stargan()
    M, S = sg.model, sg.solver
    from munch import Munch  # the real package or the shim's stand-in
    torch.set_num_threads(8)
    torch.manual_seed(777)  # main.py:227
    c = CFG
    G = M.Generator(c["img_size"], c["style_dim"], max_conv_dim=c["max_conv_dim"], w_hpf=0, norm_type="adain")
    Fm = M.MappingNetwork(c["latent_dim"], c["style_dim"], c["num_domains"])
    E = M.StyleEncoder(c["img_size"], c["style_dim"], c["num_domains"], max_conv_dim=c["max_conv_dim"])
    D = M.Discriminator(c["img_size"], c["num_domains"], max_conv_dim=c["max_conv_dim"])
    # he_init of the reference (core/utils.py he_init: kaiming_normal_ for Conv2d / Linear, bias 0)
    for net in (G, Fm, E, D):
        for m in net.modules():
            if isinstance(m, (torch.nn.Conv2d, torch.nn.Linear)):
                torch.nn.init.kaiming_normal_(m.weight, mode="fan_in", nonlinearity="relu")
                if m.bias is not None:
                    torch.nn.init.constant_(m.bias, 0)
    # non-trivial affine IN parameters so that their gradients are exercised
    gen = torch.Generator().manual_seed(1)
    for n_, p in G.named_parameters():
        if "norm" in n_ and p.dim() == 1 and "fc" not in n_ or n_.startswith("to_rgb.0"):
            p.data.add_(0.1 * torch.randn(p.shape, generator=gen))
    from oracle import starganv2_ref as R
    f_state = R.mapping_state(4242, c["latent_dim"], c["style_dim"], c["num_domains"])
    assert list(f_state) == list(Fm.state_dict())
    Fm.load_state_dict(f_state)
    nets = Munch(generator=G, mapping_network=Fm, style_encoder=E, discriminator=D)
    states = {k: {n: v.detach().clone() for n, v in net.state_dict().items()} for k, net in nets.items() if k != "mapping_network"}
    f_check = float(sum(v.double().sum() for v in f_state.values()))
    args = Munch(norm_type="adain", num_embeds=1, w_hpf=0, DiffAugment="", lambda_reg=c["lambda_reg"], lambda_sty=c["lambda_sty"],
                 lambda_ds=c["lambda_ds"], lambda_cyc=c["lambda_cyc"])
    B = c["batch"]
    x_real = torch.randn(B, 3, c["img_size"], c["img_size"], generator=gen).clamp_(-1, 1)
    x_ref = torch.randn(B, 3, c["img_size"], c["img_size"], generator=gen).clamp_(-1, 1)
    x_ref2 = torch.randn(B, 3, c["img_size"], c["img_size"], generator=gen).clamp_(-1, 1)
    y_org = torch.tensor([0, 2])
    y_trg = torch.tensor([1, 0])
    z_trg = torch.randn(B, c["latent_dim"], generator=gen)
    z_trg2 = torch.randn(B, c["latent_dim"], generator=gen)

    with torch.no_grad():
        s_map = Fm(z_trg, y_trg)
        s_enc = E(x_ref, y_trg)
        x_fake = G(x_real, s_map)
        d_out = D(x_real, y_org)
    fwd = dict(s_map=s_map, s_enc=s_enc, x_fake=x_fake, d_out=d_out)

    def grads(net):
        return {n: p.grad.detach().clone() for n, p in net.named_parameters() if p.grad is not None}

    def zero():
        for net in nets.values():
            net.zero_grad(set_to_none=True)

    res = {}
    zero()
    loss, log = S.compute_d_loss(nets, args, x_real.clone(), y_org, y_trg, z_trg=z_trg)
    loss.backward()
    res["d_latent"] = dict(loss=loss.item(), log=dict(log), grads=grads(D))
    zero()
    loss, log = S.compute_d_loss(nets, args, x_real.clone(), y_org, y_trg, x_ref=x_ref)
    loss.backward()
    res["d_ref"] = dict(loss=loss.item(), log=dict(log))
    zero()
    loss, log = S.compute_g_loss(nets, args, x_real.clone(), y_org, y_trg, z_trgs=[z_trg, z_trg2])
    loss.backward()
    res["g_latent"] = dict(loss=loss.item(), log=dict(log), grads_g=grads(G), grads_f={n: g_ for n, g_ in grads(Fm).items() if g_.numel() <= 20000}, grads_e=grads(E))
    zero()
    loss, log = S.compute_g_loss(nets, args, x_real.clone(), y_org, y_trg, x_refs=[x_ref, x_ref2])
    loss.backward()
    res["g_ref"] = dict(loss=loss.item(), log=dict(log))
    os.makedirs(OUT, exist_ok=True)
    path = os.path.join(OUT, "starganv2_small.pt")
    torch.save(dict(cfg=c, states=states, mapping_seed=4242, mapping_checksum=f_check, x_real=x_real, x_ref=x_ref, x_ref2=x_ref2, y_org=y_org, y_trg=y_trg, z_trg=z_trg,
                    z_trg2=z_trg2, fwd=fwd, res=res), path)
    print("starganv2_small.pt", os.path.getsize(path) // 1024, "KiB")
    for k, v in res.items():
        print(k, v["loss"], v["log"])


if __name__ == "__main__":
    main()
