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

if __name__ == "__main__":
    cases = [(32, 256, 66, 256, 3, 1, 0), (32, 128, 64, 512, 3, 1, 1), (32, 128, 256, 256, 3, 1, 1), (32, 128, 258, 64, 3, 1, 0)]
    knobs = [{}, {"tma_dbg": 1}, {"tma_dbg": 3}]
    run(cases, knobs)
