"""TEST INFRASTRUCTURE ONLY - generates tests/golden/*.pt by running the UNMODIFIED reference
(/root/reference/defectGAN, imported through oracle/ref_shim.py) on CPU in FP32.

    python -m oracle.make_golden            # from the repo root, in the build container

The reference cannot travel to the GPU box, so these fixtures are what pins both the oracle restatement
(tests/test_oracle_golden.py) and the CUDA path (tests/test_gpu_golden.py) to the reference's own outputs.
Networks are built with reduced widths (ngf=8, ndf=4, hidden_nc=16) so the files stay small; the code path
exercised (module classes, forward, losses, autograd) is exactly the reference's.
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_shim  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _capture_noise(ref, store):
    """Wrap NoiseInjection.forward (architecture.py:385-389) to record the drawn noise; the arithmetic
    stays the reference's own (we call the original forward with the explicit noise argument)."""
    orig = ref.architecture.NoiseInjection.forward

    def fwd(self, image, noise=None):
        if noise is None:
            b, _, h, w = image.shape
            noise = image.new_empty(b, 1, h, w).normal_()
        store.append(noise.detach().clone())
        return orig(self, image, noise)

    ref.architecture.NoiseInjection.forward = fwd
    return orig


def main():
    ref = ref_shim.import_reference()
    from utils.util import fix_rand_seed  # type: ignore  (reference: utils/util.py:21-32)
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(8)

    # ---------------- generator: forward + backward, train and eval mode ----------------
    fix_rand_seed(123)
    opt = ref_shim.default_opt(image_size=128, ngf=8, ndf=4, hidden_nc=16)
    G = ref.DefectGanGenerator(opt)
    D = ref.DefectGanDiscriminator(opt)
    G.init_weights("normal", 0.02)
    D.init_weights("normal", 0.02)
    for n, p in G.named_parameters():
        if "noise" in n:
            p.data.fill_(0.1)  # they init to 0 (architecture.py:378); exercise the noise path
    g0 = {k: v.detach().clone() for k, v in G.state_dict().items()}
    d0 = {k: v.detach().clone() for k, v in D.state_dict().items()}

    gen = torch.Generator().manual_seed(0)
    B, H = 2, 64
    x = torch.randn(B, 3, H, H, generator=gen).clamp_(-1, 1)
    labels = (torch.rand(B, 6, generator=gen) > 0.5).float()
    labels[:, 0] = 0
    noises = []
    orig = _capture_noise(ref, noises)
    G.train()
    out, prob = G(x, labels.reshape(B, 6, 1, 1))
    loss = out.square().mean() + prob.mean()
    loss.backward()
    g_grads = {n: p.grad.detach().clone() for n, p in G.named_parameters() if p.grad is not None}
    g1 = {k: v.detach().clone() for k, v in G.state_dict().items()
          if k.endswith(("weight_u", "weight_v", "running_mean", "running_var"))}
    G.eval()
    noises_eval = []
    ref.architecture.NoiseInjection.forward = orig
    _capture_noise(ref, noises_eval)
    with torch.no_grad():
        out_e, prob_e = G(x, labels.reshape(B, 6, 1, 1))
    # spatial label map (inference path, models/defectgan_model.py:385-392, test_defectgan.py:219-230)
    seg = torch.zeros(B, 6, 8, 8)
    seg[:, 0] = 1
    seg[:, 0, 2:5, 3:7] = 0
    seg[:, 3, 2:5, 3:7] = 1
    noises_seg = []
    ref.architecture.NoiseInjection.forward = orig
    _capture_noise(ref, noises_seg)
    with torch.no_grad():
        out_s, prob_s = G(x, seg)
    ref.architecture.NoiseInjection.forward = orig
    torch.save(dict(opt=vars(opt), state=g0, x=x, labels=labels, noises=noises, out=out.detach(), prob=prob.detach(),
                    grads=g_grads, buffers_after=g1, noises_eval=noises_eval, out_eval=out_e, prob_eval=prob_e,
                    seg=seg, noises_seg=noises_seg, out_seg=out_s, prob_seg=prob_s),
               os.path.join(OUT, "generator_small.pt"))

    # ---------------- discriminator: forward + backward (train: power iteration) ----------------
    HD = 128
    xd = torch.randn(B, 3, HD, HD, generator=gen).clamp_(-1, 1)
    D.train()
    src, cls = D(xd)
    (src.mean() + cls.square().mean()).backward()
    d_grads = {n: p.grad.detach().clone() for n, p in D.named_parameters()}
    d1 = {k: v.detach().clone() for k, v in D.state_dict().items() if k.endswith(("weight_u", "weight_v"))}
    D.eval()
    with torch.no_grad():
        src_e, cls_e = D(xd)
    torch.save(dict(opt=vars(opt), state=d0, x=xd, src=src.detach(), cls=cls.detach(), grads=d_grads,
                    buffers_after=d1, src_eval=src_e, cls_eval=cls_e),
               os.path.join(OUT, "discriminator_small.pt"))

    # ---------------- one D-step + one G-step through the reference's own DefectGanModel ----------------
    fix_rand_seed(123)
    from models.defectgan_model import DefectGanModel  # type: ignore
    opt = ref_shim.default_opt(image_size=128, ngf=8, ndf=4, hidden_nc=16)
    opt.device = torch.device("cpu")
    model = DefectGanModel(opt)
    model.init_weights()
    for n, p in model.netG.named_parameters():
        if "noise" in n:
            p.data.fill_(0.05)
    g0 = {k: v.detach().clone() for k, v in model.netG.state_dict().items()}
    d0 = {k: v.detach().clone() for k, v in model.netD.state_dict().items()}
    optD = torch.optim.Adam(model.netD.parameters(), lr=2e-4, betas=(0.5, 0.999))  # base_trainer.py:75-89
    optG = torch.optim.Adam(model.netG.parameters(), lr=2e-4, betas=(0.5, 0.999))
    bg = torch.randn(B, 3, HD, HD, generator=gen).clamp_(-1, 1)
    df = torch.randn(B, 3, HD, HD, generator=gen).clamp_(-1, 1)
    lab = (torch.rand(B, 6, generator=gen) > 0.5).float()
    lab[:, 0] = 0
    noises_d, noises_g = [], []
    steps = []
    w = dict(clf_d=2, clf_g=5, rec=5, sd_cyc=5, sd_con=1)  # defectgan_options.py:100-101
    for it in range(2):
        nd, ng = [], []
        _capture_noise(ref, nd)
        optD.zero_grad()  # defectgan_trainer.py:170-180
        gan_d, clf_d = model("discriminator", bg, lab, df)
        (gan_d + clf_d * w["clf_d"]).backward()
        optD.step()
        ref.architecture.NoiseInjection.forward = orig
        _capture_noise(ref, ng)
        optG.zero_grad()  # defectgan_trainer.py:138-168
        gan, clf, rec, cyc, con = model("generator", bg, lab, df)
        (gan + clf * w["clf_g"] + rec * w["rec"] + cyc * w["sd_cyc"] + con * w["sd_con"]).backward()
        if it == 0:
            g_step_grads = {n: p.grad.detach().clone() for n, p in model.netG.named_parameters() if p.grad is not None}
        optG.step()
        ref.architecture.NoiseInjection.forward = orig
        noises_d.append(nd)
        noises_g.append(ng)
        steps.append(dict(gan_d=gan_d.item(), clf_d=clf_d.item(), gan_g=gan.item(), clf_g=clf.item(),
                          rec=rec.item(), cyc=cyc.item(), con=con.item()))
    g2 = {k: v.detach().clone() for k, v in model.netG.state_dict().items()}
    d2 = {k: v.detach().clone() for k, v in model.netD.state_dict().items()}
    torch.save(dict(opt={k: v for k, v in vars(opt).items() if k != "device"}, g_state=g0, d_state=d0, bg=bg, df=df,
                    labels=lab, noises_d=noises_d, noises_g=noises_g, steps=steps, g_step_grads=g_step_grads,
                    g_state_after=g2, d_state_after=d2),
               os.path.join(OUT, "train_step_small.pt"))
    for f in sorted(os.listdir(OUT)):
        print(f, os.path.getsize(os.path.join(OUT, f)) // 1024, "KiB")
    print(steps)


if __name__ == "__main__":
    main()
