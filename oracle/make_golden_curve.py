"""TEST INFRASTRUCTURE ONLY - 100-step loss curves of the UNMODIFIED reference (DefectGanModel + torch.optim.Adam, the two
step functions of trainers/defectgan_trainer.py:138-180) on CPU in FP32 -> tests/golden/loss_curve_100.pt.

    python -m oracle.make_golden_curve       # from the repo root, in the build container

Starts from the states of tests/golden/train_step_small.pt (reduced widths: ngf 8, ndf 4, hidden_nc 16; 128x128, batch 2).
Inputs cycle through 4 seeded batches; NoiseInjection draws are replaced on both sides by the closed-form ``curve_noise``
below (CPU and CUDA generators cannot agree), indexed by a running call counter.  North-star: "loss curves tracking over
100 steps".
"""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_shim  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
STEPS = 100


def curve_noise(call: int, n: int, h: int, w: int) -> torch.Tensor:
    """Deterministic stand-in for N(0,1) noise maps: unit-variance, different for every call."""
    i = torch.arange(n * h * w, dtype=torch.float32)
    return (1.41421356 * torch.sin(i * 0.37 + 0.61 * call + 0.05 * h)).reshape(n, 1, h, w)


def curve_batches(image_size=128, batch=2, count=4, seed=77):
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(count):
        bg = torch.randn(batch, 3, image_size, image_size, generator=g).clamp_(-1, 1)
        df = torch.randn(batch, 3, image_size, image_size, generator=g).clamp_(-1, 1)
        lab = (torch.rand(batch, 6, generator=g) > 0.5).float()
        lab[:, 0] = 0
        out.append((bg, lab, df))
    return out


def main():
    ref = ref_shim.import_reference()
    from models.defectgan_model import DefectGanModel  # type: ignore
    torch.set_num_threads(8)
    t = torch.load(os.path.join(OUT, "train_step_small.pt"), weights_only=False)
    opt = ref_shim.default_opt(**{k: v for k, v in t["opt"].items()})
    opt.device = torch.device("cpu")
    model = DefectGanModel(opt)
    model.netG.load_state_dict(t["g_state"])
    model.netD.load_state_dict(t["d_state"])
    counter = [0]
    orig = ref.architecture.NoiseInjection.forward

    def fwd(self, image, noise=None):
        b, _, h, w = image.shape
        noise = curve_noise(counter[0], b, h, w)
        counter[0] += 1
        return orig(self, image, noise)

    ref.architecture.NoiseInjection.forward = fwd
    optD = torch.optim.Adam(model.netD.parameters(), lr=2e-4, betas=(0.5, 0.999))
    optG = torch.optim.Adam(model.netG.parameters(), lr=2e-4, betas=(0.5, 0.999))
    w = dict(clf_d=2, clf_g=5, rec=5, sd_cyc=5, sd_con=1)
    batches = curve_batches()
    curve = []
    for it in range(STEPS):
        bg, lab, df = batches[it % len(batches)]
        optD.zero_grad()
        gan_d, clf_d = model("discriminator", bg, lab, df)
        (gan_d + clf_d * w["clf_d"]).backward()
        optD.step()
        optG.zero_grad()
        gan, clf, rec, cyc, con = model("generator", bg, lab, df)
        (gan + clf * w["clf_g"] + rec * w["rec"] + cyc * w["sd_cyc"] + con * w["sd_con"]).backward()
        optG.step()
        curve.append([gan_d.item(), clf_d.item(), gan.item(), clf.item(), rec.item(), cyc.item(), con.item()])
        if it % 10 == 0:
            print(it, [round(v, 4) for v in curve[-1]], flush=True)
    ref.architecture.NoiseInjection.forward = orig
    torch.save(dict(names=["gan_d", "clf_d", "gan_g", "clf_g", "rec", "cyc", "con"], curve=torch.tensor(curve), steps=STEPS,
                    noise_calls=counter[0]), os.path.join(OUT, "loss_curve_100.pt"))
    print("noise calls", counter[0])


if __name__ == "__main__":
    main()
