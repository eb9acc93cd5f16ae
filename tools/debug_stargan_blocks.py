"""Debug aid: per-block (teacher-forced) error of the mirrored StarGAN-v2 generator against the FP32 CPU oracle."""
import os, sys, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import starganv2_ref as R
from de_i2i_gan_b200.stargan_v2.core import model as M
from de_i2i_gan_b200 import functional as Fn
t = torch.load("tests/golden/starganv2_small.pt", weights_only=False)
c = t["cfg"]
G = M.Generator(c["img_size"], c["style_dim"], max_conv_dim=c["max_conv_dim"], w_hpf=0).cuda()
G.load_state_dict(t["states"]["generator"])
st = t["states"]["generator"]
x, s = t["x_real"], t["fwd"]["s_map"]
rel = lambda a, b: ((a.float().cpu() - b).abs().max() / b.abs().max()).item()
def dev(h): return Fn.to_nhwc(h.cuda())
def host(a): return Fn.from_nhwc(a).cpu()
with torch.no_grad():
    h = R._conv(st, "from_rgb.", x, 1)
    print("from_rgb", rel(host(Fn.conv_linear(dev(x), G.from_rgb)), h))
    n = len(G.encode)
    for i in range(n):
        h2 = R.res_blk(st, f"encode.{i}.", h, downsample=i < n - 2)
        print("encode", i, tuple(h2.shape), rel(host(G.encode[i](dev(h))), h2))
        h = h2
    for i in range(n):
        h2 = R.adain_res_blk(st, f"decode.{i}.", h, s, upsample=i >= 2)
        print("decode", i, tuple(h2.shape), rel(host(G.decode[i](dev(h), s.cuda())), h2))
        h = h2
    print("full", rel(G(x.cuda(), s.cuda()), t["fwd"]["x_fake"]))
