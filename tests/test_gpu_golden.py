"""The CUDA path (mirrored nn.Modules -> autograd Functions -> libdei2i.so) against fixtures produced by the UNMODIFIED
reference on CPU in FP32 (oracle/make_golden.py; tests/golden/*.pt).  Same state dict, same inputs, same injected noise.

Tolerances (north-star): BF16 storage / tensor-core path vs the FP32 reference: 2e-2 relative (max |a-b| / max |b|) for
activations, outputs and losses; whole-network gradients by global relative L2 (8e-2) and per-parameter cosine (0.9),
see _grad_check; buffers updated in FP32 (u, v, running stats) 1e-2.  Per-layer gradients at 2e-2 are the business of
tests/test_gpu_kernels.py (same inputs to one layer).
"""
import json
import os

import pytest
import torch

from conftest import rel_err, GOLDEN, ROOT

pytestmark = pytest.mark.gpu

TOL_ACT = 2e-2
TOL_BUF = 1e-2


def _load(name):
    return torch.load(os.path.join(GOLDEN, name), weights_only=False)


def _report(name, **kv):
    try:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "golden_parity.jsonl"), "a") as f:
            f.write(json.dumps(dict(test=name, **kv)) + "\n")
    except OSError:
        pass


def _opt(d):
    from types import SimpleNamespace
    return SimpleNamespace(**d)


def _queue_noise(noises):
    from de_i2i_gan_b200.models.networks import architecture as A
    q = [n.clone() for n in noises]

    def src(n, h, w, device):
        t = q.pop(0)
        assert tuple(t.shape) == (n, 1, h, w), (t.shape, (n, 1, h, w))
        return t
    A.set_noise_source(src)
    return q


def _clear_noise():
    from de_i2i_gan_b200.models.networks import architecture as A
    A.set_noise_source(None)


def _grad_check(named_params, ref_grads, tag, l2_bar=8e-2):
    """Whole-network gradients after a ~30-layer BF16 backward chain: rounding compounds through the normalisation
    layers' cancellations, so element-wise max-norm bounds are meaningless for the small-gradient parameters (the CPU
    oracle with BF16 rounding emulated at the same points shows the same: tests/test_bf16_emulation.py).  Checked:
    global relative L2 error < 8e-2 (emulated BF16 level: 2.4e-2 for G, 4.7e-2 for D) and cosine > 0.9 for every parameter carrying > 1e-3 of the gradient norm."""
    got, ref = {}, {}
    for n, p in named_params:
        if n not in ref_grads:
            assert p.grad is None or p.grad.abs().max().item() == 0.0, f"{n}: reference has no gradient here"
            continue
        assert p.grad is not None, n
        got[n], ref[n] = p.grad.detach().float().cpu().flatten(), ref_grads[n].flatten()
    a, b = torch.cat(list(got.values())), torch.cat(list(ref.values()))
    rel_l2 = ((a - b).norm() / b.norm()).item()
    assert rel_l2 < l2_bar, (tag, "global rel L2", rel_l2)
    gnorm = b.norm().item()
    worst = 1.0
    for n in ref:
        if ref[n].norm().item() < 1e-3 * gnorm:
            continue
        cos = (torch.dot(got[n], ref[n]) / (got[n].norm() * ref[n].norm() + 1e-30)).item()
        worst = min(worst, cos)
        assert cos > 0.9, (tag, n, cos)
    return dict(rel_l2=rel_l2, worst_cos=worst)


def test_generator_train_forward_backward():
    from de_i2i_gan_b200.models.networks.generator import DefectGanGenerator
    g = _load("generator_small.pt")
    G = DefectGanGenerator(_opt(g["opt"]))
    G.load_state_dict(g["state"])
    G.cuda().train()
    q = _queue_noise(g["noises"])
    try:
        x = g["x"].cuda()
        labels = g["labels"].cuda()
        out, prob = G(x, labels.reshape(labels.shape[0], -1, 1, 1))
        assert not q, "noise draws out of order / count"
    finally:
        _clear_noise()
    e_out, e_prob = rel_err(out, g["out"]), rel_err(prob, g["prob"])
    (out.square().mean() + prob.mean()).backward()
    worst = _grad_check(G.named_parameters(), g["grads"], "G")
    bufs = G.state_dict()
    e_buf = max(rel_err(bufs[k], v) for k, v in g["buffers_after"].items())
    _report("golden/generator_train", out=e_out, prob=e_prob, grads=worst, buffers=e_buf)
    assert e_out < TOL_ACT and e_prob < TOL_ACT, (e_out, e_prob)
    assert e_buf < TOL_BUF, e_buf


def test_generator_eval_and_spatial_labels():
    from de_i2i_gan_b200.models.networks.generator import DefectGanGenerator
    g = _load("generator_small.pt")
    G = DefectGanGenerator(_opt(g["opt"]))
    st = dict(g["state"])
    st.update(g["buffers_after"])
    G.load_state_dict(st)
    G.cuda().eval()
    x = g["x"].cuda()
    try:
        with torch.no_grad():
            _queue_noise(g["noises_eval"])
            out, prob = G(x, g["labels"].cuda().reshape(x.shape[0], -1, 1, 1))
            _queue_noise(g["noises_seg"])
            out_s, prob_s = G(x, g["seg"].cuda())
    finally:
        _clear_noise()
    errs = dict(out=rel_err(out, g["out_eval"]), prob=rel_err(prob, g["prob_eval"]), out_seg=rel_err(out_s, g["out_seg"]),
                prob_seg=rel_err(prob_s, g["prob_seg"]))
    _report("golden/generator_eval", **errs)
    assert max(errs.values()) < TOL_ACT, errs
    # eval mode must not touch the spectral-norm / batch-norm state
    after = G.state_dict()
    for k, v in g["buffers_after"].items():
        assert torch.equal(after[k].cpu(), v), k


def test_discriminator_forward_backward():
    from de_i2i_gan_b200.models.networks.discriminator import DefectGanDiscriminator
    d = _load("discriminator_small.pt")
    D = DefectGanDiscriminator(_opt(d["opt"]))
    D.load_state_dict(d["state"])
    D.cuda().train()
    x = d["x"].cuda()
    src, cls = D(x)
    e_src, e_cls = rel_err(src, d["src"]), rel_err(cls, d["cls"])
    (src.mean() + cls.square().mean()).backward()
    worst = _grad_check(D.named_parameters(), d["grads"], "D")
    bufs = D.state_dict()
    e_buf = max(rel_err(bufs[k], v) for k, v in d["buffers_after"].items())
    D.eval()
    with torch.no_grad():
        src_e, cls_e = D(x)
    e_eval = max(rel_err(src_e, d["src_eval"]), rel_err(cls_e, d["cls_eval"]))
    _report("golden/discriminator", src=e_src, cls=e_cls, grads=worst, buffers=e_buf, eval=e_eval)
    assert e_src < TOL_ACT and e_cls < TOL_ACT and e_eval < TOL_ACT, (e_src, e_cls, e_eval)
    assert e_buf < TOL_BUF, e_buf


def test_two_training_steps_track_reference():
    """Two D-step + G-step iterations through DefectGanModel + FlatAdam (the trainer's body) vs the reference's own
    DefectGanModel + torch.optim.Adam: every loss within 2e-2, first G-step gradients, and the direction of the
    two-step parameter update (Adam normalises element-wise, so noise-level gradients flip signs: cosine > 0.9)."""
    from de_i2i_gan_b200.models.defectgan_model import DefectGanModel
    from de_i2i_gan_b200.optim import FlatAdam
    t = _load("train_step_small.pt")
    opt = _opt(t["opt"])
    opt.device = torch.device("cuda", 0)
    model = DefectGanModel(opt)
    model.netG.load_state_dict(t["g_state"])
    model.netD.load_state_dict(t["d_state"])
    model.netG.cuda()
    model.netD.cuda()
    optD = FlatAdam(list(model.netD.parameters()), lr=2e-4, betas=(0.5, 0.999))
    optG = FlatAdam(list(model.netG.parameters()), lr=2e-4, betas=(0.5, 0.999))
    w = dict(clf_d=2, clf_g=5, rec=5, sd_cyc=5, sd_con=1)
    bg, df, lab = t["bg"].cuda(), t["df"].cuda(), t["labels"].cuda()
    worst_loss = 0.0
    try:
        for it in range(2):
            q = _queue_noise(t["noises_d"][it])
            optD.zero_grad()
            gan_d, clf_d = model("discriminator", bg, lab, df)
            (gan_d + clf_d * w["clf_d"]).backward()
            optD.step()
            assert not q
            q = _queue_noise(t["noises_g"][it])
            optG.zero_grad()
            gan, clf, rec, cyc, con = model("generator", bg, lab, df)
            (gan + clf * w["clf_g"] + rec * w["rec"] + cyc * w["sd_cyc"] + con * w["sd_con"]).backward()
            assert not q
            if it == 0:
                worst_grad = _grad_check(model.netG.named_parameters(), t["g_step_grads"], "G-step")
            optG.step()
            got = dict(gan_d=gan_d.item(), clf_d=clf_d.item(), gan_g=gan.item(), clf_g=clf.item(), rec=rec.item(),
                       cyc=cyc.item(), con=con.item())
            for k, v in t["steps"][it].items():
                err = abs(got[k] - v) / max(abs(v), 1e-3)
                worst_loss = max(worst_loss, err)
                assert err < TOL_ACT, (it, k, got[k], v)
    finally:
        _clear_noise()
    cos = {}
    for name, net in (("g", model.netG), ("d", model.netD)):
        before, after = t[name + "_state"], t[name + "_state_after"]
        keys = [n for n, _ in net.named_parameters()]
        st = net.state_dict()
        d_ref = torch.cat([(after[k] - before[k]).flatten() for k in keys])
        d_got = torch.cat([(st[k].detach().cpu() - before[k]).flatten() for k in keys])
        cos[name] = (torch.dot(d_ref, d_got) / (d_ref.norm() * d_got.norm())).item()
    _report("golden/train_two_steps", worst_loss_rel=worst_loss, grads=worst_grad, cos_g=cos["g"], cos_d=cos["d"])
    assert cos["g"] > 0.9 and cos["d"] > 0.9, cos
