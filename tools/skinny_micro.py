"""Micro-benchmark of the skinny convolution geometries of the C2 workload through the C ABI (CUDA-event timing)."""
import sys, os
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from de_i2i_gan_b200 import ops
from conv_micro import bench

def case(n, cin, h, cout, k, s, pad, seg1x1=False, which="fdw", iters=5):
    spec = ops.ConvSpec(cin, cout, k, s, pad, "zeros")
    hs = 1 if seg1x1 else h
    x = torch.randn(n, hs, hs, spec.cin_ld, device="cuda").bfloat16()
    w = torch.randn(cout, cin, k, k, device="cuda") * 0.02
    wf, wd = ops.pack_weight(w, spec)
    ho, wo = spec.out_hw(h, h)
    dy = torch.randn(n, ho, wo, spec.cout_ld, device="cuda").bfloat16()
    dw = torch.empty_like(w)
    fl = spec.flops(n, ho, wo)
    tag = spec.tag(n, h, h) + (" seg1x1" if seg1x1 else "")
    lhw = (h, h) if seg1x1 else None
    out_b = n * ho * wo * spec.cout_ld * 2
    if "f" in which:
        ms = bench(f"fwd   {tag}", lambda: ops.conv2d_fwd(x, wf, spec, logical_hw=lhw), fl, iters)
        print(f"      output {out_b/1e6:.0f} MB -> {out_b/ms/1e6:.0f} GB/s")
    if "d" in which and not seg1x1:
        bench(f"dgrad {tag}", lambda: ops.conv2d_dgrad(dy, wd, spec, (h, h)), fl, iters)
    if "w" in which:
        ms = bench(f"wgrad {tag}", lambda: ops.conv2d_wgrad(x, dy, dw, spec, logical_hw=lhw), fl, iters)
        print(f"      dy {out_b/1e6:.0f} MB -> {out_b/ms/1e6:.0f} GB/s")

if __name__ == "__main__":
    sel = sys.argv[1] if len(sys.argv) > 1 else "all"
    if sel in ("all", "spade"):
        case(32, 6, 256, 128, 3, 1, 1, True, "fw")
        case(32, 6, 64, 128, 3, 1, 1, True, "fw")
    if sel in ("all", "stem"):
        case(32, 3, 262, 64, 7, 1, 0, False, "fdw")
    if sel in ("all", "heads"):
        case(32, 64, 258, 4, 3, 1, 0, False, "fdw")
    if sel in ("all", "d"):
        case(32, 3, 258, 64, 4, 2, 0, False, "fw")
        case(32, 2048, 6, 1, 3, 1, 0, False, "fdw")
        case(32, 2048, 4, 6, 4, 1, 0, False, "fdw")
        case(32, 1024, 10, 2048, 4, 2, 0, False, "fdw")
