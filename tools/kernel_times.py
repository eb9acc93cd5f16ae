#!/usr/bin/env python
"""Per-kernel device time of ONE eager D-step + G-step (torch.profiler / CUPTI, no kernel replay: seconds instead of the
minutes of an ncu launch list).  Guidance for optimisation only - bench numbers come from bench.py, evidence from ncu."""
import os
import sys
from collections import defaultdict
from pathlib import Path

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402


def main():
    from de_i2i_gan_b200 import options
    from de_i2i_gan_b200.trainers import find_trainer_using_model_name
    from de_i2i_gan_b200 import synthetic as R
    B, H = int(os.environ.get("B", 32)), int(os.environ.get("H", 256))
    torch.manual_seed(123)
    opt = options.readme_recipe(image_size=H, batch_size=B, num_critics=1, num_epochs=10, iters_per_epoch=10 ** 6, name="kt",
                                ckpt_dir=Path("/tmp/dei2i_kt"))
    opt.cuda_graph = False
    tr = find_trainer_using_model_name(opt.model)(opt)
    bg, lab, df = [t.cuda() for t in R.synthetic_batch(B, H, seed=0)]
    for _ in range(3):
        tr._train_discriminator_once(bg, lab, df)
        tr._train_generator_once(bg, lab, df)
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as prof:
        tr._train_discriminator_once(bg, lab, df)
        tr._train_generator_once(bg, lab, df)
        torch.cuda.synchronize()
    agg = defaultdict(lambda: [0, 0.0])
    for e in prof.events():
        if e.device_type == torch.autograd.DeviceType.CUDA and e.device_time > 0:
            name = e.name.split("(")[0][:90]
            agg[name][0] += 1
            agg[name][1] += e.device_time
    total = sum(v[1] for v in agg.values())
    print(f"kernels {sum(v[0] for v in agg.values())}  device time {total / 1e3:.2f} ms")
    for name, (n, us) in sorted(agg.items(), key=lambda kv: -kv[1][1])[:45]:
        print(f"{us / 1e3:8.3f} ms {100 * us / total:5.1f}%  x{n:5d}  avg {us / n:8.1f} us  {name}")


if __name__ == "__main__":
    main()
