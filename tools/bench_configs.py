#!/usr/bin/env python
"""Device-timed throughput of the other BASELINE.json configs on one B200 (bench.py stays the configs[1] headline):

  configs[2]  WGAN (train_wgan.py) weight-clipping step, 128x128, batch 64          -> img/s
  configs[3]  StarGAN-v2 iteration (2 D-losses with R1 + 2 G-losses + EMA), 256x256, batch 16 -> img/s
  configs[4]  DefectGAN inference-only defect synthesis (_generate_fake), 256x256, batch 1..256 -> img/s per batch size

One JSON line per measurement (CUDA events around K repetitions after W warm-ups; synthetic inputs resident in HBM).
    python tools/bench_configs.py [--which wgan,stargan,inference] [--steps K] [--warmup W]
"""
import argparse
import json
import os
import sys
from pathlib import Path
from types import SimpleNamespace

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

import torch  # noqa: E402


def timed(fn, steps, warmup):
    for _ in range(warmup):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / steps


def emit(**kv):
    print(json.dumps(kv), flush=True)


def bench_wgan(args):
    from de_i2i_gan_b200 import options
    from de_i2i_gan_b200.trainers.wgan_trainer import WGanTrainer
    B, H = 64, 128
    opt = options.wgan_options(image_size=H, batch_size=B, ckpt_dir=Path("/tmp/dei2i_wgan_bench"), num_display_images=4)
    torch.manual_seed(123)
    tr = WGanTrainer(opt)
    real = torch.randn(B, 3, H, H, device="cuda").clamp_(-1, 1)

    def step():
        tr._train_discriminator_once(real)
        tr._train_generator_once(B)
    ms = timed(step, args.steps, args.warmup)
    emit(config="configs[2] WGAN weight-clipping D-step + G-step, 128x128, batch 64", ms_per_step=ms, img_per_s=B / ms * 1e3,
         peak_mem_gb=round(torch.cuda.max_memory_allocated() / 2 ** 30, 2))


def bench_stargan(args):
    from de_i2i_gan_b200 import functional as Fn
    from de_i2i_gan_b200.optim import FlatAdam
    from de_i2i_gan_b200.stargan_v2.core import model as M, solver as S
    B, H = 16, 256
    a = SimpleNamespace(img_size=H, style_dim=64, latent_dim=16, num_domains=3, w_hpf=0, norm_type="adain", num_embeds=1,
                        lambda_reg=1.0, lambda_sty=1.0, lambda_ds=2.0, lambda_cyc=1.0)
    torch.manual_seed(777)
    nets, nets_ema = M.build_nets(a)
    for d in (nets, nets_ema):
        for k, v in vars(d).items():
            for m in v.modules():
                if isinstance(m, (torch.nn.Conv2d, torch.nn.Linear)):  # he_init, core/utils.py
                    torch.nn.init.kaiming_normal_(m.weight, mode="fan_in", nonlinearity="relu")
            v.cuda()
    # solver.py:72-80: Adam(lr 1e-4 (mapping network 1e-6), betas (0, 0.99), weight_decay 1e-4)
    opts = {k: FlatAdam(list(v.parameters()), lr=1e-6 if k == "mapping_network" else 1e-4, betas=(0.0, 0.99), weight_decay=1e-4)
            for k, v in vars(nets).items()}
    x = torch.randn(B, 3, H, H, device="cuda").clamp_(-1, 1)
    xr, xr2 = torch.randn_like(x).clamp_(-1, 1), torch.randn_like(x).clamp_(-1, 1)
    y_org = torch.randint(0, 3, (B,), device="cuda")
    y_trg = torch.randint(0, 3, (B,), device="cuda")
    z, z2 = torch.randn(B, 16, device="cuda"), torch.randn(B, 16, device="cuda")

    def zero():
        for o in opts.values():
            o.zero_grad()

    def iteration():  # solver.py:258-308
        with Fn.weight_cache():
            zero(); loss, _ = S.compute_d_loss(nets, a, x.clone(), y_org, y_trg, z_trg=z); loss.backward()
        opts["discriminator"].step()
        with Fn.weight_cache():
            zero(); loss, _ = S.compute_d_loss(nets, a, x.clone(), y_org, y_trg, x_ref=xr); loss.backward()
        opts["discriminator"].step()
        with Fn.weight_cache():
            zero(); loss, _ = S.compute_g_loss(nets, a, x, y_org, y_trg, z_trgs=[z, z2]); loss.backward()
        opts["generator"].step(); opts["mapping_network"].step(); opts["style_encoder"].step()
        with Fn.weight_cache():
            zero(); loss, _ = S.compute_g_loss(nets, a, x, y_org, y_trg, x_refs=[xr, xr2]); loss.backward()
        opts["generator"].step()
        S.moving_average(nets.generator, nets_ema.generator, beta=0.999)
        S.moving_average(nets.mapping_network, nets_ema.mapping_network, beta=0.999)
        S.moving_average(nets.style_encoder, nets_ema.style_encoder, beta=0.999)
    ms = timed(iteration, args.steps, args.warmup)
    emit(config="configs[3] StarGAN-v2 iteration (2x D-loss with R1 double backward, 2x G-loss, EMA), 256x256, batch 16",
         ms_per_iteration=ms, img_per_s=B / ms * 1e3, peak_mem_gb=round(torch.cuda.max_memory_allocated() / 2 ** 30, 2))


def bench_inference(args):
    from de_i2i_gan_b200 import options
    from de_i2i_gan_b200.models import create_model
    H = 256
    opt = options.readme_recipe(image_size=H, batch_size=1, ckpt_dir=Path("/tmp/dei2i_inf_bench"), phase="test", is_train=False)
    torch.manual_seed(123)
    model = create_model(opt)
    model.init_weights()
    for b in (1, 4, 16, 64, 256):
        x = torch.randn(b, 3, H, H, device="cuda").clamp_(-1, 1)
        lab = torch.zeros(b, 6, device="cuda")
        lab[:, 1 + torch.arange(b) % 5] = 1.0

        def fwd():
            with torch.no_grad():
                model("inference", x, lab)
        ms = timed(fwd, max(args.steps, 5), args.warmup)
        emit(config=f"configs[4] DefectGAN inference (_generate_fake) 256x256, batch {b}", batch=b, ms_per_batch=ms, img_per_s=b / ms * 1e3)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--which", default="wgan,stargan,inference")
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    args = ap.parse_args()
    assert torch.cuda.is_available(), "needs a CUDA device (no CPU fallback on the product path)"
    for w in args.which.split(","):
        {"wgan": bench_wgan, "stargan": bench_stargan, "inference": bench_inference}[w](args)


if __name__ == "__main__":
    main()
