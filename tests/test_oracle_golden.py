"""The oracle restatement (oracle/defectgan_ref.py) against fixtures produced by the UNMODIFIED reference
(oracle/make_golden.py).  CPU only; FP32; tolerance 1e-5 relative (same math, different op grouping)."""
import os

import torch

from oracle import defectgan_ref as R
from conftest import rel_err, GOLDEN

TOL = 1e-5


def _load(name):
    return torch.load(os.path.join(GOLDEN, name), weights_only=False)


def test_generator_forward_backward_matches_reference():
    g = _load("generator_small.pt")
    st = R.clone_state(g["state"])
    for k in R.trainable_keys(st):
        st[k].requires_grad_(True)
    out, prob = R.generator_forward(st, g["x"], g["labels"], True, R.NoiseSource(g["noises"]))
    assert rel_err(out, g["out"]) < TOL and rel_err(prob, g["prob"]) < TOL
    (out.square().mean() + prob.mean()).backward()
    gmax = max(v.abs().max().item() for v in g["grads"].values())
    for n, ref in g["grads"].items():
        assert (st[n].grad - ref).abs().max().item() < 1e-4 * ref.abs().max().item() + 1e-7 * gmax, n
    dead = [k for k in R.trainable_keys(st) if st[k].grad is None]
    assert sorted(dead) == sorted(k for k in R.trainable_keys(st) if k not in g["grads"])  # norm_s/conv_s never run
    for k, ref in g["buffers_after"].items():
        assert rel_err(st[k], ref) < TOL, k


def test_generator_eval_and_spatial_labels():
    g = _load("generator_small.pt")
    st = R.clone_state(g["state"])
    st.update(R.clone_state(g["buffers_after"]))
    with torch.no_grad():
        out, prob = R.generator_forward(st, g["x"], g["labels"], False, R.NoiseSource(g["noises_eval"]))
        assert rel_err(out, g["out_eval"]) < TOL and rel_err(prob, g["prob_eval"]) < TOL
        out, prob = R.generator_forward(st, g["x"], g["seg"], False, R.NoiseSource(g["noises_seg"]))
        assert rel_err(out, g["out_seg"]) < TOL and rel_err(prob, g["prob_seg"]) < TOL


def test_discriminator_matches_reference():
    d = _load("discriminator_small.pt")
    st = R.clone_state(d["state"])
    for k in R.trainable_keys(st):
        st[k].requires_grad_(True)
    src, cls = R.discriminator_forward(st, d["x"], True)
    assert rel_err(src, d["src"]) < TOL and rel_err(cls, d["cls"]) < TOL
    (src.mean() + cls.square().mean()).backward()
    for n, ref in d["grads"].items():
        assert rel_err(st[n].grad, ref) < 1e-4, n
    for k, ref in d["buffers_after"].items():
        assert rel_err(st[k], ref) < TOL, k
    with torch.no_grad():
        src, cls = R.discriminator_forward(st, d["x"], False)
    assert rel_err(src, d["src_eval"]) < TOL and rel_err(cls, d["cls_eval"]) < TOL


def test_two_training_steps_match_reference():
    t = _load("train_step_small.pt")
    tr = R.Trainer(R.clone_state(t["g_state"]), R.clone_state(t["d_state"]))
    for it in range(2):
        gan_d, clf_d = tr.train_discriminator_once(t["bg"], t["labels"], t["df"], R.NoiseSource(t["noises_d"][it]))
        losses = tr.train_generator_once(t["bg"], t["labels"], t["df"], R.NoiseSource(t["noises_g"][it]))
        ref = t["steps"][it]
        got = dict(gan_d=gan_d, clf_d=clf_d, gan_g=losses[0], clf_g=losses[1], rec=losses[2], cyc=losses[3], con=losses[4])
        for k, v in ref.items():
            assert abs(got[k] - v) < 1e-4 * max(abs(v), 1e-3), (it, k, got[k], v)
        if it == 0:  # gradients of the first G-step, element-wise
            gmax = max(v.abs().max().item() for v in t["g_step_grads"].values())
            for n, ref_g in t["g_step_grads"].items():
                assert (tr.g[n].grad - ref_g).abs().max().item() < 5e-3 * ref_g.abs().max().item() + 1e-5 * gmax, n
    # Adam normalises each element's gradient, so an element whose gradient is rounding noise moves by
    # +-lr with a noise-determined sign: compare the two-step update of each whole network by cosine.
    for name, after in (("g", t["g_state_after"]), ("d", t["d_state_after"])):
        st = tr.g if name == "g" else tr.d
        before = t[name + "_state"]
        keys = [k for k in R.trainable_keys(after)]
        d_ref = torch.cat([(after[k] - before[k]).flatten() for k in keys])
        d_got = torch.cat([(st[k].detach() - before[k]).flatten() for k in keys])
        cos = torch.dot(d_ref, d_got) / (d_ref.norm() * d_got.norm())
        assert cos > 0.99, (name, cos.item())
