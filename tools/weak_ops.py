#!/usr/bin/env python
"""Runs the weakest kernels of the C2 step in isolation (for `ncu --set full -k regex:...` captures and quick CUDA-event timing).
   python tools/weak_ops.py [name ...]   names: fwd64 wgrad64 stem d5 spade_bwd bn heads"""
import os
import sys

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch  # noqa: E402
from de_i2i_gan_b200 import ops  # noqa: E402


def timeit(tag, fn, reps=5):
    for _ in range(2):
        fn()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    torch.cuda.synchronize()
    e0.record()
    for _ in range(reps):
        fn()
    e1.record()
    torch.cuda.synchronize()
    print(f"{tag:60s} {e0.elapsed_time(e1) / reps * 1e3:9.1f} us   [{ops.last_kernel()}]", flush=True)


def conv_case(n, cin, h, w, cout, kh, kw, s, which="fdw"):
    spec = ops.ConvSpec(cin, cout, kh, s, 0, "zeros", kw if kw != kh else 0)
    x = torch.randn(n, h, w, spec.cin_ld, device="cuda").bfloat16()
    wt = torch.randn(cout, cin, kh, kw, device="cuda") * 0.02
    wf, wd = ops.pack_weight(wt, spec)
    ho, wo = spec.out_hw(h, w)
    dy = torch.randn(n, ho, wo, spec.cout_ld, device="cuda").bfloat16()
    dw = torch.empty_like(wt)
    tag = spec.tag(n, h, w)
    fl = spec.flops(n, ho, wo)
    if "f" in which:
        timeit(f"fwd   {tag} ({fl / 1e9:.0f} GF)", lambda: ops.conv2d_fwd(x, wf, spec))
    if "d" in which:
        timeit(f"dgrad {tag}", lambda: ops.conv2d_dgrad(dy, wd, spec, (h, w)))
    if "w" in which:
        timeit(f"wgrad {tag}", lambda: ops.conv2d_wgrad(x, dy, dw, spec))


def main():
    if os.environ.get("MMA_ISSUE"):
        ops.set_tuning("mma_issue", int(os.environ["MMA_ISSUE"]))
    if os.environ.get("TMA_PAIR"):
        ops.set_tuning("tma_pair", int(os.environ["TMA_PAIR"]))
    names = sys.argv[1:] or ["fwd64", "wgrad64", "stem", "d5", "spade_bwd", "bn", "heads"]
    g = torch.Generator(device="cuda").manual_seed(0)
    if "fwd64" in names:
        conv_case(32, 128, 258, 258, 64, 3, 3, 1, "fd")
    if "wgrad64" in names:
        conv_case(32, 128, 258, 258, 64, 3, 3, 1, "w")
        ops.set_tuning("wgrad_swap", 0)
        conv_case(32, 128, 258, 258, 64, 3, 3, 1, "w")
        ops.set_tuning("wgrad_swap", 1)
    if "n128" in names:
        conv_case(32, 256, 130, 130, 128, 3, 3, 1, "fd")
        conv_case(32, 64, 258, 258, 128, 4, 4, 2, "f")
    if "c256" in names:
        conv_case(32, 256, 66, 66, 256, 3, 3, 1, "fdw")
    if "stem" in names:
        conv_case(32, 64, 262, 256, 64, 7, 1, 1, "fdw")
        ops.set_tuning("wgrad_ring", 0)
        conv_case(32, 64, 262, 256, 64, 7, 1, 1, "w")
        ops.set_tuning("wgrad_ring", 1)
    if "d5" in names:
        conv_case(32, 1024, 10, 10, 2048, 4, 4, 2, "fdw")
        conv_case(32, 512, 18, 18, 1024, 4, 4, 2, "fdw")
    if "dlayers" in names:  # the discriminator's strided 4x4 chain (pre-padded inputs) at 256^2 and the generator's two encoder convs
        for cin, hin in ((64, 130), (128, 66), (256, 34), (512, 18), (1024, 10)):
            conv_case(32, cin, hin, hin, 2 * cin, 4, 4, 2, "d")
        conv_case(32, 64, 258, 258, 128, 4, 4, 2, "d")
        conv_case(32, 128, 130, 130, 256, 4, 4, 2, "d")
    if "heads" in names:
        conv_case(32, 64, 258, 258, 4, 3, 3, 1, "fdw")
    if "spade_bwd" in names:
        for c, hx, up in ((128, 128, 2), (256, 64, 2), (256, 64, 1)):
            x = (torch.randn(32, hx, hx, c, generator=g, device="cuda") * 2 + 1).bfloat16()
            gb5 = (torch.randn(32, 5, 5, 2 * c, generator=g, device="cuda") * 0.5).bfloat16()
            out, mr = ops.spade_cls_fwd(x, gb5, up, out_pad=1)
            dout = torch.randn_like(out)
            timeit(f"spade_cls_fwd c{c} hx{hx} up{up}", lambda: ops.spade_cls_fwd(x, gb5, up, out_pad=1))
            timeit(f"spade_cls_bwd c{c} hx{hx} up{up}", lambda: ops.spade_cls_bwd(x, gb5, out, dout, mr, up, out_pad=1))
    if "bn" in names:
        for c, h in ((64, 256), (256, 64)):
            x = torch.randn(32, h, h, c, generator=g, device="cuda").bfloat16()
            gamma, beta = torch.ones(c, device="cuda"), torch.zeros(c, device="cuda")
            rm, rv = torch.zeros(c, device="cuda"), torch.ones(c, device="cuda")
            y, save = ops.bn_fwd(x, gamma, beta, rm, rv, True, act="leaky_relu", out_pad=1)
            dy = torch.randn_like(x)
            timeit(f"bn_fwd_pad c{c} h{h}", lambda: ops.bn_fwd(x, gamma, beta, rm, rv, True, act="leaky_relu", out_pad=1))
            timeit(f"bn_bwd c{c} h{h}", lambda: ops.bn_bwd(x, dy, gamma, beta, save, True, act="leaky_relu"))
            dyp = torch.randn_like(y)
            timeit(f"bn_bwd c{c} h{h} dy_pad=1", lambda: ops.bn_bwd(x, dyp, gamma, beta, save, True, act="leaky_relu", dy_pad=1))
            timeit(f"reflect_fold c{c} h{h}", lambda: ops.reflect_fold(y, 1))
            timeit(f"channel_stats c{c} h{h}", lambda: ops.channel_stats(x, False))


if __name__ == "__main__":
    main()
