"""TEST INFRASTRUCTURE ONLY - generates tests/golden/wgan_small.pt by running the UNMODIFIED reference networks
(WGanGenerator / WGanDiscriminator, /root/reference/defectGAN/models/networks/{generator,discriminator}.py, imported through
oracle/ref_shim.py) on CPU in FP32.

    python -m oracle.make_golden_wgan      # from the repo root, in the build container

``WGanTrainer`` itself cannot be constructed as checked in, so the two step functions
(trainers/wgan_trainer.py:115-138, models/wgan_model.py:18-21) are applied here verbatim around the reference's own
networks with torch.optim.RMSprop (base_trainer.py:72-73), with the uniform latent (generator.py:46) drawn up front.
Reduced widths (ngf = ndf = 8, 32x32 -> num_layers = 2) keep the fixture small.
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_shim  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def main():
    ref = ref_shim.import_reference()
    from utils.util import fix_rand_seed  # type: ignore
    torch.set_num_threads(8)
    fix_rand_seed(123)
    size, ngf, ndf, nz, B = 32, 8, 8, 100, 4
    L = size.bit_length() - 1 - 3
    G = ref.WGanGenerator(num_layers=L, ngf=ngf, noise_dim=nz)
    D = ref.WGanDiscriminator(num_layers=L, ndf=ndf)
    G.init_weights("normal", 0.02)
    D.init_weights("normal", 0.02)
    g0 = {k: v.detach().clone() for k, v in G.state_dict().items()}
    d0 = {k: v.detach().clone() for k, v in D.state_dict().items()}
    gen = torch.Generator().manual_seed(5)
    z = torch.rand(B, nz, 1, 1, generator=gen)
    real = torch.randn(B, 3, size, size, generator=gen).clamp_(-1, 1)

    # forward + backward of each network (train mode)
    G.train(); D.train()
    fake = G(z)
    fake.square().mean().backward()
    g_grads = {n: p.grad.detach().clone() for n, p in G.named_parameters()}
    xin = real.clone().requires_grad_(True)
    logits = D(xin)
    logits.sum().backward()
    d_grads = {n: p.grad.detach().clone() for n, p in D.named_parameters()}
    dx = xin.grad.detach().clone()
    G.eval(); D.eval()
    with torch.no_grad():
        fake_eval = G(z)
        logits_eval = D(real)
    fwd = dict(fake=fake.detach(), g_grads=g_grads, logits=logits.detach(), d_grads=d_grads, dx=dx, fake_eval=fake_eval,
               logits_eval=logits_eval)

    # two D-steps + G-steps from the initial state
    G.load_state_dict(g0); D.load_state_dict(d0)
    G.train(); D.train()
    lr, clip = 5e-5, 0.03  # options/wgan_options.py:52,57
    optG = torch.optim.RMSprop(G.parameters(), lr=lr)
    optD = torch.optim.RMSprop(D.parameters(), lr=lr)
    zs = [torch.rand(B, nz, 1, 1, generator=gen) for _ in range(4)]
    steps = []
    for it in range(2):
        for p in D.parameters():  # wgan_model.py:18-21
            p.data.clamp_(-clip, clip)
        optD.zero_grad()  # wgan_trainer.py:124-138
        fake_data = G(zs[2 * it])
        fake_logits = D(fake_data.detach_())
        real_logits = D(real)
        d_loss = -(real_logits.mean() - fake_logits.mean())
        d_loss.backward()
        optD.step()
        optG.zero_grad()  # wgan_trainer.py:115-122
        fake_data = G(zs[2 * it + 1])
        g_loss = -D(fake_data).mean()
        g_loss.backward()
        optG.step()
        steps.append(dict(d_loss=d_loss.item(), g_loss=g_loss.item()))
    g2 = {k: v.detach().clone() for k, v in G.state_dict().items()}
    d2 = {k: v.detach().clone() for k, v in D.state_dict().items()}
    os.makedirs(OUT, exist_ok=True)
    torch.save(dict(cfg=dict(image_size=size, ngf=ngf, ndf=ndf, noise_dim=nz, batch=B, lr=lr, clipping_limit=clip),
                    g_state=g0, d_state=d0, z=z, real=real, fwd=fwd, zs=zs, steps=steps, g_state_after=g2, d_state_after=d2),
               os.path.join(OUT, "wgan_small.pt"))
    print("wgan_small.pt", os.path.getsize(os.path.join(OUT, "wgan_small.pt")) // 1024, "KiB", steps)


if __name__ == "__main__":
    main()
