"""One small invocation of the hot path on cuda:0, checked against the CPU oracle (used by __graft_entry__.smoke()).

Reduced-width DefectGAN generator + discriminator (ngf=16, ndf=8, hidden_nc=32) at 64x64 / 128x128, batch 2: one
generator forward + backward and one discriminator forward + backward through the mirrored modules (-> libdei2i.so),
compared with oracle/defectgan_ref.py on the same state dict, inputs and injected noise.  BF16 path vs FP32 oracle:
2e-2 relative on outputs."""
import torch


def run(verbose=True):
    from types import SimpleNamespace
    from oracle import defectgan_ref as R  # checker only (test infrastructure)
    from . import ops
    from .models.networks import architecture as A
    from .models.networks.generator import DefectGanGenerator
    from .models.networks.discriminator import DefectGanDiscriminator

    dev = torch.device("cuda", 0)
    opt = SimpleNamespace(num_res=6, cycle_gan=False, label_nc=6, skip_conn=False, ngf=16, input_nc=3, use_spectral=True,
                          num_scales=2, add_noise=True, style_norm_block_type="spade", hidden_nc=32, style_distill=False,
                          embed_nc=768, ndf=8, num_layers=5, image_size=128)
    g_state = R.random_generator_state(ngf=16, hidden_nc=32, noise_weight=0.1)
    d_state = R.random_discriminator_state(ndf=8, image_size=128)
    bg, labels, _ = R.synthetic_batch(2, 64, seed=0)
    xd, _, _ = R.synthetic_batch(2, 128, seed=1)

    def rel(a, b):
        a, b = a.detach().float().cpu(), b.detach().float().cpu()
        return ((a - b).abs().max() / b.abs().max().clamp_min(1e-20)).item()

    def rel_l2(params, ref_state):
        """global relative L2 error of all gradients (BF16 backward chain vs FP32 oracle)"""
        keys = [k for k in R.trainable_keys(ref_state) if ref_state[k].grad is not None]
        a = torch.cat([params[k].grad.detach().float().cpu().flatten() for k in keys])
        b = torch.cat([ref_state[k].grad.flatten() for k in keys])
        return ((a - b).norm() / b.norm()).item()

    launches0 = ops.launch_count()
    # ---- generator
    G = DefectGanGenerator(opt)
    G.load_state_dict(g_state)
    G.to(dev).train()
    gen = torch.Generator().manual_seed(7)
    shapes = [(2, 1, 16, 16)] * 6 + [(2, 1, 32, 32), (2, 1, 64, 64)]
    noises = [torch.randn(s, generator=gen) for s in shapes]
    q = [n.clone() for n in noises]
    A.set_noise_source(lambda n, h, w, device: q.pop(0))
    try:
        out, prob = G(bg.to(dev), labels.to(dev))
        (out.square().mean() + prob.mean()).backward()
    finally:
        A.set_noise_source(None)
    st = R.clone_state(g_state)
    for k in R.trainable_keys(st):
        st[k].requires_grad_(True)
    o_ref, p_ref = R.generator_forward(st, bg, labels, True, R.NoiseSource(noises))
    (o_ref.square().mean() + p_ref.mean()).backward()
    e = dict(g_out=rel(out, o_ref), g_prob=rel(prob, p_ref))
    e["g_grad"] = rel_l2(dict(G.named_parameters()), st)
    # ---- discriminator
    D = DefectGanDiscriminator(opt)
    D.load_state_dict(d_state)
    D.to(dev).train()
    src, cls = D(xd.to(dev))
    (src.mean() + cls.square().mean()).backward()
    sd = R.clone_state(d_state)
    for k2 in R.trainable_keys(sd):
        sd[k2].requires_grad_(True)
    s_ref, c_ref = R.discriminator_forward(sd, xd, True)
    (s_ref.mean() + c_ref.square().mean()).backward()
    e["d_src"], e["d_cls"] = rel(src, s_ref), rel(cls, c_ref)
    e["d_grad"] = rel_l2(dict(D.named_parameters()), sd)
    torch.cuda.synchronize()
    e["launches"] = ops.launch_count() - launches0
    if verbose:
        print("smoke:", {k: (round(v, 5) if isinstance(v, float) else v) for k, v in e.items()})
    assert e["launches"] > 0, "no CUDA kernels of libdei2i.so were launched"
    for name in ("g_out", "g_prob", "d_src", "d_cls"):
        assert e[name] < 2e-2, (name, e[name])
    for name in ("g_grad", "d_grad"):
        assert e[name] < 8e-2, (name, e[name])  # BF16 storage alone gives 2.4e-2 (G) / 4.7e-2 (D): tests/test_bf16_emulation.py
    return e
