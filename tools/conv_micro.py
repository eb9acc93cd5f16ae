"""Micro-benchmark of single convolution geometries through the C ABI (CUDA-event timing, inputs >> L2 rotate)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from de_i2i_gan_b200 import ops

def bench(name, fn, flops, iters=10):
    for _ in range(3): fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters): fn()
    e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    print(f"{name:60s} {ms*1e3:9.1f} us  {flops/ms/1e9:8.1f} TF/s", flush=True)
    return ms

def run(cases, knobs):
    for (n, cin, h, cout, k, s, pad) in cases:
        spec = ops.ConvSpec(cin, cout, k, s, pad, "zeros")
        x = torch.randn(n, h, h, spec.cin_ld, device="cuda").bfloat16()
        w = torch.randn(cout, cin, k, k, device="cuda") * 0.02
        wf, wd = ops.pack_weight(w, spec)
        ho, wo = spec.out_hw(h, h)
        dy = torch.randn(n, ho, wo, spec.cout_ld, device="cuda").bfloat16()
        dw = torch.empty_like(w)
        fl = spec.flops(n, ho, wo)
        tag = spec.tag(n, h, h)
        for kn in knobs:
            for key, val in kn.items(): ops.set_tuning(key, val)
            bench(f"fwd   {tag} {kn}", lambda: ops.conv2d_fwd(x, wf, spec), fl)
            if s == 1:
                bench(f"dgrad {tag} {kn}", lambda: ops.conv2d_dgrad(dy, wd, spec, (h, h)), fl)
            for key in kn: ops.set_tuning(key, 0)
        bench(f"wgrad {tag}", lambda: ops.conv2d_wgrad(x, dy, dw, spec), fl)

def run_epilogues():
    """forward 256->256 3x3 with the epilogue variants the generator uses (1/sigma scale, noise, residual)"""
    n, cin, h, cout, k = 32, 256, 66, 256, 3
    spec = ops.ConvSpec(cin, cout, k, 1, 0, "zeros")
    x = torch.randn(n, h, h, spec.cin_ld, device="cuda").bfloat16()
    w = torch.randn(cout, cin, k, k, device="cuda") * 0.02
    wf, _ = ops.pack_weight(w, spec, need_dgrad=False)
    ho, wo = spec.out_hw(h, h)
    fl = spec.flops(n, ho, wo)
    noise = torch.randn(n, 1, ho, wo, device="cuda")
    nw, sc = torch.tensor([0.1], device="cuda"), torch.tensor([0.7], device="cuda")
    res = torch.randn(n, ho, wo, cout, device="cuda").bfloat16()
    bench("fwd plain", lambda: ops.conv2d_fwd(x, wf, spec), fl, 20)
    bench("fwd scale", lambda: ops.conv2d_fwd(x, wf, spec, scale=sc), fl, 20)
    bench("fwd scale+noise", lambda: ops.conv2d_fwd(x, wf, spec, scale=sc, noise=noise, noise_weight=nw), fl, 20)
    bench("fwd scale+noise+residual", lambda: ops.conv2d_fwd(x, wf, spec, scale=sc, noise=noise, noise_weight=nw, residual=res), fl, 20)
    bench("fwd scale+lrelu", lambda: ops.conv2d_fwd(x, wf, spec, scale=sc, act="leaky_relu"), fl, 20)
    # sustained: 300 back-to-back launches (power cap)
    bench("fwd plain x300 (sustained)", lambda: ops.conv2d_fwd(x, wf, spec), fl, 300)
    bench("fwd scale+noise+residual x300", lambda: ops.conv2d_fwd(x, wf, spec, scale=sc, noise=noise, noise_weight=nw, residual=res), fl, 300)


def run_spade_gb():
    """SPADE gamma|beta convolutions (K = 1152, bias): the epilogue-heavy dense geometry"""
    for (n, cin, h, cout) in [(32, 128, 64, 512), (32, 128, 256, 256)]:
        spec = ops.ConvSpec(cin, cout, 3, 1, 1, "zeros")
        x = torch.randn(n, h, h, spec.cin_ld, device="cuda").bfloat16()
        w = torch.randn(cout, cin, 3, 3, device="cuda") * 0.02
        wf, _ = ops.pack_weight(w, spec, need_dgrad=False)
        fl = spec.flops(n, h, h)
        bias = torch.randn(cout, device="cuda")
        tag = spec.tag(n, h, h)
        bench(f"fwd plain {tag}", lambda: ops.conv2d_fwd(x, wf, spec), fl, 20)
        bench(f"fwd bias  {tag}", lambda: ops.conv2d_fwd(x, wf, spec, bias=bias), fl, 20)
        for dbg in (1, 3):
            ops.set_tuning("tma_dbg", dbg)
            bench(f"fwd bias dbg={dbg} (1: no stores, 3: no stores no tmem loads) {tag}", lambda: ops.conv2d_fwd(x, wf, spec, bias=bias), fl, 20)
            ops.set_tuning("tma_dbg", 0)


if __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "gb":
    run_spade_gb()
elif __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "epi":
    run_epilogues()
elif __name__ == "__main__" and len(sys.argv) > 1 and sys.argv[1] == "prof":
    # the two dominant dense geometries, default knobs only (ncu --set full target)
    run([(32, 256, 66, 256, 3, 1, 0), (32, 128, 64, 512, 3, 1, 1)], [{}])
elif __name__ == "__main__":
    cases = [(32, 256, 66, 256, 3, 1, 0), (32, 128, 64, 512, 3, 1, 1), (32, 128, 256, 256, 3, 1, 1), (32, 128, 258, 64, 3, 1, 0)]
    knobs = [{}, {"tma_dbg": 1}, {"tma_dbg": 3}]
    run(cases, knobs)
