"""Debug aid: gradient error of the pieces of encode.0 (ResBlk 64->16, affine IN, down-sampling) vs FP32 torch on CPU."""
import os, sys, torch, torch.nn.functional as F
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import starganv2_ref as R
from de_i2i_gan_b200.stargan_v2.core import model as M
from de_i2i_gan_b200 import functional as Fn
t = torch.load("tests/golden/starganv2_small.pt", weights_only=False)
c = t["cfg"]
G = M.Generator(c["img_size"], c["style_dim"], max_conv_dim=c["max_conv_dim"], w_hpf=0).cuda()
G.load_state_dict(t["states"]["generator"])
st = {k: v.clone() for k, v in t["states"]["generator"].items()}
blk = G.encode[0]
rel = lambda a, b: ((a.float().cpu() - b).abs().max() / b.abs().max()).item()
l2 = lambda a, b: ((a.float().cpu() - b).norm() / b.norm()).item()
with torch.no_grad():
    h = R._conv(st, "from_rgb.", t["x_real"], 1).to(torch.bfloat16).float()
gen = torch.Generator().manual_seed(9)
p = "encode.0."
def run(name, ref_fn, dev_fn):
    hin = h.clone().requires_grad_(True)
    ref = ref_fn(hin)
    dout = torch.randn(ref.shape, generator=gen).to(torch.bfloat16).float()
    (dref,) = torch.autograd.grad(ref, hin, dout)
    xin = h.clone().cuda().requires_grad_(True)
    with Fn.weight_cache():
        out = Fn.from_nhwc(dev_fn(Fn.to_nhwc(xin)))
        (dgot,) = torch.autograd.grad(out, xin, dout.cuda())
    print(f"{name:28s} out {rel(out, ref):.4f}  dx max {rel(dgot, dref):.4f}  dx l2 {l2(dgot, dref):.4f}")
n1 = lambda x: F.leaky_relu(R._in_affine(st, p + "norm1.", x), 0.2)
run("norm1+lrelu", n1, lambda x: blk._norm_act(x, blk.norm1))
run("norm1+lrelu+conv1", lambda x: R._conv(st, p + "conv1.", n1(x), 1), lambda x: Fn.conv_linear(blk._norm_act(x, blk.norm1), blk.conv1))
run("conv1 only", lambda x: R._conv(st, p + "conv1.", x, 1), lambda x: Fn.conv_linear(x, blk.conv1))
run("conv1x1 only", lambda x: R._conv(st, p + "conv1x1.", x, 0), lambda x: Fn.conv_linear(x, blk.conv1x1))
run("avgpool", lambda x: F.avg_pool2d(x, 2), lambda x: Fn.avg_pool(x, 2))
run("shortcut", lambda x: F.avg_pool2d(R._conv(st, p + "conv1x1.", x, 0), 2), lambda x: blk._shortcut(x))
run("residual", lambda x: R.res_blk(st, p, x, True) * 2 ** 0.5 - F.avg_pool2d(R._conv(st, p + "conv1x1.", x, 0), 2), lambda x: blk._residual(x))
run("block", lambda x: R.res_blk(st, p, x, True), lambda x: blk(x))
