"""TEST INFRASTRUCTURE ONLY - tests/golden/mae_small.pt from the UNMODIFIED reference (DefectGanModel in the MAE pre-training
modes, defectGAN/models/defectgan_model.py:106-171,361-383; MaskToken, architecture.py:392-418) on CPU in FP32.

    python -m oracle.make_golden_mae

Reduced-width nets (ngf=8, ndf=4, hidden_nc=16), 128x128, batch 2; for each mask-token type the fixture holds the losses of
'mae_discriminator' and 'mae_generator', the generator / mask-token gradients of the G loss, the mask that the seeded host RNG
produced and the injected noise - what tests/test_gpu_mae.py replays through the CUDA path."""
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_shim  # noqa: E402
from oracle.make_golden import _capture_noise  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def main():
    ref = ref_shim.import_reference()
    from utils.util import fix_rand_seed  # type: ignore
    from models.defectgan_model import DefectGanModel  # type: ignore
    torch.set_num_threads(8)
    B, H = 2, 128
    gen = torch.Generator().manual_seed(0)
    imgs = torch.randn(B, 3, H, H, generator=gen).clamp_(-1, 1)
    labels = (torch.rand(B, 6, generator=gen) > 0.5).float()
    out = dict(imgs=imgs, labels=labels, cases={})
    for ttype in ("position", "zero", "mean", "vector", "scalar", "full"):
        fix_rand_seed(123)
        opt = ref_shim.default_opt(image_size=H, ngf=8, ndf=4, hidden_nc=16, mask_token_type=ttype, mask_ratio=0.75, patch_size=8,
                                   split_training=False)
        opt.device = torch.device("cpu")
        model = DefectGanModel(opt)
        model.init_weights()
        for n, p in model.netG.named_parameters():
            if "noise" in n:
                p.data.fill_(0.05)
        if isinstance(model.mask_token.mask_token, torch.nn.Parameter):
            with torch.no_grad():
                model.mask_token.mask_token.copy_(0.3 * torch.randn(model.mask_token.mask_token.shape, generator=gen))
        case = dict(opt={k: v for k, v in vars(opt).items() if k != "device"},
                    g_state={k: v.detach().clone() for k, v in model.netG.state_dict().items()},
                    d_state={k: v.detach().clone() for k, v in model.netD.state_dict().items()},
                    token=(model.mask_token.mask_token.detach().clone() if isinstance(model.mask_token.mask_token, torch.nn.Parameter) else None))
        # D loss
        nd = []
        orig = _capture_noise(ref, nd)
        torch.manual_seed(77)
        gan_d, clf_d = model("mae_discriminator", imgs, labels)
        (gan_d + 3 * clf_d).backward()
        case["d"] = dict(gan=gan_d.item(), clf=clf_d.item(), noises=nd,
                         grads={n: p.grad.detach().clone() for n, p in model.netD.named_parameters() if p.grad is not None})
        ref.architecture.NoiseInjection.forward = orig
        # G loss
        for p in model.netD.parameters():
            p.grad = None
        ng = []
        _capture_noise(ref, ng)
        torch.manual_seed(78)
        rec, gan, clf = model("mae_generator", imgs, labels)
        (gan + 10 * rec + clf).backward()
        ref.architecture.NoiseInjection.forward = orig
        case["g"] = dict(rec=rec.item(), gan=gan.item(), clf=clf.item(), noises=ng,
                         grads={n: p.grad.detach().clone() for n, p in model.netG.named_parameters() if p.grad is not None},
                         token_grad=(model.mask_token.mask_token.grad.detach().clone()
                                     if isinstance(model.mask_token.mask_token, torch.nn.Parameter) and model.mask_token.mask_token.grad is not None else None))
        # the masks those two seeds produce (utils/util.py:60-71), for the oracle-side pin
        from utils.util import generate_shifted_mask  # type: ignore
        torch.manual_seed(77)
        case["d"]["mask"] = generate_shifted_mask(imgs.size(), 8, 0.75).bool()
        torch.manual_seed(78)
        case["g"]["mask"] = generate_shifted_mask(imgs.size(), 8, 0.75).bool()
        out["cases"][ttype] = case
        print(ttype, case["d"]["gan"], case["d"]["clf"], case["g"]["rec"], case["g"]["gan"], case["g"]["clf"])
    # the six cases share images / labels; network states differ only through the RNG stream -> keep the file small: store the
    # states once where they are equal
    first = out["cases"]["position"]
    for t, c in out["cases"].items():
        if t != "position" and all(torch.equal(c["g_state"][k], first["g_state"][k]) for k in first["g_state"]):
            c["g_state"] = None
        if t != "position" and all(torch.equal(c["d_state"][k], first["d_state"][k]) for k in first["d_state"]):
            c["d_state"] = None
        if t != "position":  # whole-network gradients once; the other token types keep losses + the token gradient
            c["d"]["grads"] = None
            c["g"]["grads"] = None
    torch.save(out, os.path.join(OUT, "mae_small.pt"))
    print(os.path.getsize(os.path.join(OUT, "mae_small.pt")) // 1024, "KiB")


if __name__ == "__main__":
    main()
