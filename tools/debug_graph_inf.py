import os, sys, pathlib, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from de_i2i_gan_b200 import options
from de_i2i_gan_b200.models import create_model
torch.manual_seed(123)
opt = options.readme_recipe(image_size=128, batch_size=2, ngf=16, ndf=8, hidden_nc=32, ckpt_dir=pathlib.Path("/tmp/x"))
opt.device = torch.device("cuda", 0)
m = create_model(opt); m.init_weights()
g = torch.Generator().manual_seed(6)
xs = [torch.randn(2, 3, 128, 128, generator=g).clamp(-1, 1).cuda() for _ in range(3)]
lab = torch.zeros(2, 6).cuda(); lab[:, 2] = 1
with torch.no_grad():
    os.environ["DEI2I_CUDA_GRAPH"] = "0"
    e1 = m("inference", xs[2], lab)[0].clone(); e2 = m("inference", xs[2], lab)[0].clone()
    print("eager vs eager", (e1 - e2).abs().max().item())
    os.environ["DEI2I_CUDA_GRAPH"] = "1"
    m("inference", xs[0], lab); m("inference", xs[1], lab)
    g1 = m("inference", xs[2], lab)[0].clone(); g2 = m("inference", xs[2], lab)[0].clone()
    print("graph vs graph", (g1 - g2).abs().max().item(), "graph vs eager", (g1 - e1).abs().max().item())
    for p in m.netG.parameters(): p.mul_(1.01)
    g3 = m("inference", xs[2], lab)[0].clone()
    os.environ["DEI2I_CUDA_GRAPH"] = "0"
    e3 = m("inference", xs[2], lab)[0].clone()
    print("after weight change: graph vs eager", (g3 - e3).abs().max().item(), "changed by", (e3 - e1).abs().max().item())
