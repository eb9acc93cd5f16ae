#!/usr/bin/env python
"""Per-kernel device times of single convolution ops (fwd / dgrad / wgrad through the C ABI) for chosen geometries
(torch.profiler; guidance only).   python tools/op_profile.py [n,cin,h,cout,k,stride,pad ...]"""
import os
import sys
from collections import defaultdict

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from torch.profiler import ProfilerActivity, profile  # noqa: E402
from de_i2i_gan_b200 import ops  # noqa: E402

DEFAULT = [(32, 1024, 10, 2048, 4, 2, 0), (32, 512, 18, 1024, 4, 2, 0), (32, 256, 34, 512, 4, 2, 0), (32, 64, 258, 4, 3, 1, 0),
           (32, 128, 258, 64, 3, 1, 0), (32, 2048, 4, 6, 4, 1, 0)]


def prof(tag, fn, reps=5):
    for _ in range(3):
        fn()
    torch.cuda.synchronize()
    with profile(activities=[ProfilerActivity.CUDA]) as p:
        for _ in range(reps):
            fn()
        torch.cuda.synchronize()
    agg = defaultdict(float)
    for e in p.events():
        if e.device_type == torch.autograd.DeviceType.CUDA and e.device_time > 0:
            agg[e.name.split("(")[0][:70]] += e.device_time / reps
    print(f"{tag:55s} total {sum(agg.values()):8.1f} us : " + ", ".join(f"{k.replace('dei2i::', '').replace('void ', '')} {v:.1f}" for k, v in
                                                                       sorted(agg.items(), key=lambda kv: -kv[1])), flush=True)


def main():
    cases = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]] or DEFAULT
    for (n, cin, h, cout, k, s, pad) in cases:
        spec = ops.ConvSpec(cin, cout, k, s, pad, "zeros")
        x = torch.randn(n, h, h, spec.cin_ld, device="cuda").bfloat16()
        w = torch.randn(cout, cin, k, k, device="cuda") * 0.02
        wf, wd = ops.pack_weight(w, spec)
        ho, wo = spec.out_hw(h, h)
        dy = torch.randn(n, ho, wo, spec.cout_ld, device="cuda").bfloat16()
        dw = torch.empty_like(w)
        tag = spec.tag(n, h, h)
        prof("fwd   " + tag, lambda: ops.conv2d_fwd(x, wf, spec))
        prof("dgrad " + tag, lambda: ops.conv2d_dgrad(dy, wd, spec, (h, h)))
        prof("wgrad " + tag, lambda: ops.conv2d_wgrad(x, dy, dw, spec))
        u = torch.randn(cout, device="cuda")
        inv = torch.ones(2, device="cuda")
        try:
            prof("wgrad+sn " + tag, lambda: ops.conv2d_wgrad(x, dy, dw, spec, sn=(w, u, u.new_zeros(cin * k * k), inv)))
        except Exception as ex:  # signature differs: ignore
            print("  (sn variant skipped:", type(ex).__name__, str(ex)[:80], ")")


if __name__ == "__main__":
    main()
