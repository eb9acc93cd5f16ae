"""North-star "loss curves tracking over 100 steps": 100 D-steps + G-steps of the CUDA trainer against the curve of the
UNMODIFIED reference (CPU, FP32; oracle/make_golden_curve.py -> tests/golden/loss_curve_100.pt), from the same initial
state, the same cycling batches and the same closed-form noise maps."""
import os
import pathlib
from types import SimpleNamespace

import pytest
import torch

from conftest import GOLDEN, ROOT

pytestmark = pytest.mark.gpu


def test_loss_curves_track_reference_over_100_steps(tmp_path):
    from oracle.make_golden_curve import curve_batches, curve_noise  # seeded input / noise factories only
    from de_i2i_gan_b200 import options
    from de_i2i_gan_b200.models.networks import architecture as A
    from de_i2i_gan_b200.trainers import find_trainer_using_model_name
    t = torch.load(os.path.join(GOLDEN, "train_step_small.pt"), weights_only=False)
    c = torch.load(os.path.join(GOLDEN, "loss_curve_100.pt"), weights_only=False)
    o = t["opt"]
    opt = options.readme_recipe(image_size=o["image_size"], batch_size=2, ngf=o["ngf"], ndf=o["ndf"], hidden_nc=o["hidden_nc"],
                                num_critics=1, num_epochs=10, iters_per_epoch=1000, name="curve", ckpt_dir=pathlib.Path(tmp_path))
    opt.device = torch.device("cuda", 0)
    opt.cuda_graph = False  # the noise factory below is host-indexed by a call counter
    tr = find_trainer_using_model_name(opt.model)(opt)
    tr.model.netG.load_state_dict(t["g_state"])
    tr.model.netD.load_state_dict(t["d_state"])
    counter = [0]

    def noise(n, h, w, device):
        m = curve_noise(counter[0], n, h, w)
        counter[0] += 1
        return m.to(device)

    batches = [tuple(x.cuda() for x in b) for b in curve_batches()]
    A.set_noise_source(noise)
    got = []
    try:
        for it in range(c["steps"]):
            bg, lab, df = batches[it % len(batches)]
            tr._train_discriminator_once(bg, lab, df)
            tr._train_generator_once(bg, lab, df)
            L = tr.losses
            got.append([L["gan"]["D"][-1], L["clf"]["D"][-1], L["gan"]["G"][-1], L["clf"]["G"][-1], L["aux"]["rec"][-1],
                        L["aux"]["cyc"][-1], L["aux"]["con"][-1]])
    finally:
        A.set_noise_source(None)
    assert counter[0] == c["noise_calls"], "noise maps must be drawn in the reference's order and number"
    got, ref = torch.tensor(got), c["curve"]
    dev = (got - ref).abs() / ref.abs().clamp_min(0.25)
    report = {n: (dev[:, i].max().item(), dev[:, i].mean().item()) for i, n in enumerate(c["names"])}
    try:
        import json
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "loss_curve_parity.json"), "w") as f:
            json.dump(dict(report=report, got=got.tolist(), ref=ref.tolist()), f)
    except OSError:
        pass
    # BF16 storage vs the FP32 reference from a common start (deviation = |a - b| / max(|b|, 0.25)).  A GAN step is one move of
    # an adversarial game: rounding differences grow step by step, so the bar widens with the horizon (measured: 7e-3 over the
    # first 20 steps, 0.13 up to step 50; afterwards single steps of the oscillating generator-adversarial loss differ by O(1)
    # while 20-step means stay within 6e-2, profiles/r01f_loss_curve_parity.json):
    assert dev[:20].max().item() < 2e-2, report          # north-star BF16 tolerance while the trajectories coincide
    assert dev[:50].max().item() < 0.25 and dev[:50].mean().item() < 3e-2, report
    win = (got[-20:].mean(0) - ref[-20:].mean(0)).abs() / ref[-20:].mean(0).abs().clamp_min(0.25)
    assert win.max().item() < 0.12, (win.tolist(), report)   # the curves still track at the end of the 100 steps
    smooth = [i for i, n in enumerate(c["names"]) if n != "gan_g"]
    assert dev[:, smooth].mean().item() < 5e-2, report
    # and the curves actually move: the reference's reconstruction loss falls by > 20 % over the 100 steps, ours with it
    assert ref[-4:, 4].mean() < 0.8 * ref[:4, 4].mean() and got[-4:, 4].mean() < 0.8 * got[:4, 4].mean()
