"""TEST INFRASTRUCTURE ONLY - generates tests/golden/style_small.pt by running the UNMODIFIED reference
(/root/reference/defectGAN, imported through oracle/ref_shim.py) on CPU in FP32: the SEAN and AdaIN style variants of DefectGAN
(README.md:26-31), one D-loss + one G-loss through the reference's own DefectGanModel per variant.

    python -m oracle.make_golden_style            # from the repo root, in the build container

Variants: adain with sean_alpha = 1 (convolutional StyleExtractor) and 0 (label | noise MLP), sean with sean_alpha = 1 (stored
embeddings, drawn with python's `random.choices`) and 0 (label code only).  Host RNG protocol, repeated by the tests:
`random.seed(s); torch.manual_seed(s)` immediately before each model call - the extractor's noise (extractor.py:85) and the embedding
choices (defectgan_model.py:408) are the first host draws of a call, so the CUDA path (whose NoiseInjection maps are injected from the
fixture) sees the same values.  `style_distill=True` is not covered: the reference raises on it (normalization.py:189-190).
"""
import itertools
import os
import random
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import ref_shim  # noqa: E402
from oracle.make_golden import _capture_noise  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")
VARIANTS = [("adain", 1), ("adain", 0), ("sean", 1), ("sean", 0)]
W = dict(clf_d=2, clf_g=5, rec=5, sd_cyc=5, sd_con=1)  # defectgan_options.py:100-101


def style_opt(kind, alpha, embed_path=None):
    return ref_shim.default_opt(image_size=64, ngf=8, ndf=4, num_layers=4, num_res=2, hidden_nc=16, style_norm_block_type=kind, sean_alpha=alpha,
                                style_distill=False, embed_nc=24, latent_dim=16, num_embeds=3, embed_path=embed_path)


def make_embeddings(seed=5):
    """label combination -> list of stored embeddings; some combinations deliberately empty (the zeros branch, :405-406)."""
    g = torch.Generator().manual_seed(seed)
    emb = {}
    for l in itertools.product([0, 1], repeat=6):
        emb[l] = [torch.randn(24, generator=g) for _ in range(4)] if (sum(l) % 2 == 0 or l[0] == 1) else []
    return emb


def main():
    ref = ref_shim.import_reference()
    from models.defectgan_model import DefectGanModel  # type: ignore
    from utils.util import fix_rand_seed  # type: ignore
    os.makedirs(OUT, exist_ok=True)
    torch.set_num_threads(8)
    B, H = 2, 64
    gen = torch.Generator().manual_seed(11)
    bg = torch.randn(B, 3, H, H, generator=gen).clamp_(-1, 1)
    df = torch.randn(B, 3, H, H, generator=gen).clamp_(-1, 1)
    lab = torch.tensor([[0., 1., 0., 1., 0., 0.], [0., 0., 1., 0., 0., 0.]])  # first row: stored embeddings exist, second: none
    emb = make_embeddings()
    emb_path = "/tmp/dei2i_style_embeddings.pt"
    torch.save(emb, emb_path)
    cases = {}
    for kind, alpha in VARIANTS:
        fix_rand_seed(123)
        opt = style_opt(kind, alpha, emb_path if (kind == "sean" and alpha != 0) else None)
        opt.device = torch.device("cpu")
        model = DefectGanModel(opt)
        model.init_weights()
        for n, p in model.netG.named_parameters():
            if "noise" in n:
                p.data.fill_(0.05)
        states = {name: {k: v.detach().clone() for k, v in net.state_dict().items()} for name, net in model.networks.items()}
        keys = {name: list(st) for name, st in states.items()}  # every state-dict key, for the mirror's key-parity check
        # SEAN's 2 x 2^label_nc running-statistics buffers per layer are zeros until update_stats() runs: keep the key, drop the tensor
        states["G"] = {k: v for k, v in states["G"].items() if not (k.rsplit(".", 1)[-1].startswith(("mean_", "std_")) and not v.any())}
        nd, ng = [], []
        orig = _capture_noise(ref, nd)
        random.seed(7)
        torch.manual_seed(7)
        gan_d, clf_d = model("discriminator", bg, lab, df)
        model.netD.zero_grad()
        (gan_d + clf_d * W["clf_d"]).backward()
        d_names = [n for n, p in model.netD.named_parameters() if p.grad is not None]
        d_grads = {n: p.grad.detach().clone() for n, p in model.netD.named_parameters() if n in (d_names[0], d_names[-1])}
        ref.architecture.NoiseInjection.forward = orig
        _capture_noise(ref, ng)
        for net in model.networks.values():
            net.zero_grad()
        random.seed(8)
        torch.manual_seed(8)
        gan, clf, rec, cyc, con = model("generator", bg, lab, df)
        (gan + clf * W["clf_g"] + rec * W["rec"] + cyc * W["sd_cyc"] + con * W["sd_con"]).backward()
        ref.architecture.NoiseInjection.forward = orig
        g_grads = {n: p.grad.detach().clone() for n, p in model.netG.named_parameters() if p.grad is not None}
        e_grads = {}
        if "E" in model.networks:
            e_grads = {n: p.grad.detach().clone() for n, p in model.netE.named_parameters() if p.grad is not None}
            if alpha == 0:  # 256-wide MLP: the first and the last layer pin the backward through all five
                keep = list(e_grads)[:2] + list(e_grads)[-2:]
                e_grads = {n: e_grads[n] for n in keep}
        # a plain generator forward in eval mode with the style inputs of the model's own helper
        random.seed(9)
        torch.manual_seed(9)
        ne = []
        _capture_noise(ref, ne)
        model.netG.eval()
        with torch.no_grad():
            out_e, prob_e = model._generate_fake(bg, lab)
        ref.architecture.NoiseInjection.forward = orig
        if kind == "sean" and alpha == 0 and all(torch.equal(v, cases["sean_1"]["states"]["D"][k]) for k, v in states["D"].items()):
            states["D"] = "sean_1"  # same seed, same architecture: stored once
        cases[f"{kind}_{alpha}"] = dict(
            opt={k: v for k, v in vars(opt).items() if k not in ("device", "embed_path")}, states=states, keys=keys, noises_d=nd, noises_g=ng,
            noises_eval=ne, losses=dict(gan_d=gan_d.item(), clf_d=clf_d.item(), gan_g=gan.item(), clf_g=clf.item(), rec=rec.item(),
                                        cyc=cyc.item(), con=con.item()),
            d_grads=d_grads, g_grads=g_grads, e_grads=e_grads, out_eval=out_e, prob_eval=prob_e)
        print(kind, alpha, cases[f"{kind}_{alpha}"]["losses"], "noise draws", len(nd), len(ng), len(ne))
    path = os.path.join(OUT, "style_small.pt")
    torch.save(dict(bg=bg, df=df, labels=lab, embeddings=emb, cases=cases), path)
    print(path, os.path.getsize(path) // 1024, "KiB")


if __name__ == "__main__":
    main()
