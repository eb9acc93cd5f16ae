"""Micro-benchmark / ncu target of the class-map SPADE kernels at the generator's geometries (batch 32)."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from de_i2i_gan_b200 import ops

def run(n, hx, c, up, pad, reps):
    x = torch.randn(n, hx, hx, c, device="cuda").to(torch.bfloat16)
    gb5 = (0.3 * torch.randn(n, 5, 5, 2 * c, device="cuda")).to(torch.bfloat16)
    out, mr = ops.spade_cls_fwd(x, gb5, up, True, 1e-5, pad)
    dout = torch.randn_like(out)
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    torch.cuda.synchronize()
    ev[0].record()
    for _ in range(reps):
        ops.spade_cls_fwd(x, gb5, up, True, 1e-5, pad)
    ev[1].record()
    for _ in range(reps):
        ops.spade_cls_bwd(x, gb5, out, dout, mr, up, True, pad)
    ev[2].record()
    torch.cuda.synchronize()
    h = hx * up
    fwd_b = (x.numel() + out.numel()) * 2
    bwd_b = 2 * (x.numel() + 2 * out.numel()) * 2 + x.numel() * 2
    tf, tb = ev[0].elapsed_time(ev[1]) / reps, ev[1].elapsed_time(ev[2]) / reps
    print(f"N{n} {hx}->{h} C{c} pad{pad}: fwd(+stats) {tf * 1e3:7.1f} us {fwd_b / tf / 1e6:7.0f} GB/s | bwd1+bwd2 {tb * 1e3:7.1f} us {bwd_b / tb / 1e6:7.0f} GB/s")

if __name__ == "__main__":
    reps = 1 if len(sys.argv) > 1 and sys.argv[1] == "prof" else 10
    for cfg in ((32, 64, 256, 1, 1), (32, 64, 256, 2, 1), (32, 128, 128, 2, 1)):
        run(*cfg, reps)
        if reps > 1:
            run(*cfg, reps)
