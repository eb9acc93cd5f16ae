#!/usr/bin/env python
"""Diagnostic: per-parameter gradient error of the G-step loss graph vs the FP32 torch oracle on the GPU."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
torch.backends.cudnn.allow_tf32 = False
torch.backends.cuda.matmul.allow_tf32 = False
import test_gpu_full_width as T
from oracle import defectgan_ref as R
from de_i2i_gan_b200.synthetic import synthetic_batch
from de_i2i_gan_b200.models.defectgan_model import DefectGanModel

def run(B, H):
    g_state, d_state = T._states(H)
    model = DefectGanModel(T._opt(H, B))
    model.netG.load_state_dict(g_state); model.netD.load_state_dict(d_state)
    model.netG.cuda(); model.netD.cuda()
    bg, labels, df = synthetic_batch(B, H, seed=2)
    noises = T._draw(T._noise_shapes(B, H, 4), 23)
    b, l, d = bg.cuda(), labels.cuda(), df.cuda()
    w = dict(zip(("clf_d", "clf_g", "rec", "sd_cyc", "sd_con"), R.LOSS_WEIGHT_DEFAULT))
    for p in model.netD.parameters(): p.requires_grad_(False)
    q = T._inject(noises)
    try:
        gan, clf, rec, cyc, con = model("generator", b, l, d)
        (gan + clf * w["clf_g"] + rec * w["rec"] + cyc * w["sd_cyc"] + con * w["sd_con"]).backward()
    finally:
        T._clear()
    dev = torch.device("cuda", 0)
    g_ref, d_ref = T._req_grad(T._to(g_state, dev)), T._to(d_state, dev)
    lr = R.generator_loss(g_ref, d_ref, b, l, d, "bce", R.NoiseSource([n.cuda() for n in noises]))
    (lr[0] + lr[1] * w["clf_g"] + lr[2] * w["rec"] + lr[3] * w["sd_cyc"] + lr[4] * w["sd_con"]).backward()
    ref = T._ref_grads(g_ref)
    # the same FP32 oracle with BF16 storage emulated at every activation / gradient: what the storage format alone does
    from oracle.bf16_emulation import bf16_storage, q as q16
    g_em = T._req_grad(T._to(g_state, dev))
    with bf16_storage():
        le = R.generator_loss(g_em, T._to(d_state, dev), q16(b), l, q16(d), "bce", R.NoiseSource([n.cuda() for n in noises]))
        (le[0] + le[1] * w["clf_g"] + le[2] * w["rec"] + le[3] * w["sd_cyc"] + le[4] * w["sd_con"]).backward()
    emu = T._ref_grads(g_em)
    rows = []
    tot_e = tot_n = 0.0
    for n, p in model.netG.named_parameters():
        if n not in ref: continue
        a, r = p.grad.detach().float().cpu().flatten(), ref[n].flatten()
        e = (a - r).norm().item(); rn = r.norm().item()
        tot_e += e * e; tot_n += rn * rn
        ee = (emu[n].flatten() - r).norm().item()
        rows.append((e, rn, n, ee))
    print(f"B={B} H={H}: global rel L2 {(tot_e / tot_n) ** 0.5:.4f}  losses ours {[round(x.item(),5) for x in (gan,clf,rec,cyc,con)]} ref {[round(x.item(),5) for x in lr]}")
    rows.sort(reverse=True)
    ge = sum(r[3] ** 2 for r in rows) ** 0.5 / tot_n ** 0.5
    print(f"   emulated-BF16 oracle vs FP32 oracle: global rel L2 {ge:.4f}")
    for e, rn, n, ee in rows[:14]:
        print(f"   ours rel {e / max(rn, 1e-30):8.4f}   emulated-bf16 rel {ee / max(rn, 1e-30):8.4f}   ref norm {rn:10.4e}  {n}")
    del model, g_ref, d_ref
    torch.cuda.empty_cache()

for B, H in [(int(a), int(b)) for a, b in (s.split("x") for s in (sys.argv[1:] or ["8x256", "4x256", "8x128"]))]:
    run(B, H)
