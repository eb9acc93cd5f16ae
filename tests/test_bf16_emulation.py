"""What BF16 storage alone costs through the whole generator (CPU, no kernels): the FP32 oracle vs the same oracle with
BF16 rounding emulated at every activation / gradient (oracle/bf16_emulation.py).  This pins the tolerances the GPU
golden tests use for whole-network gradients: outputs stay within 2e-2, the global gradient within 5e-2 relative L2,
while single small-gradient parameters lose 20-30 % in max norm (cancellation in BatchNorm / InstanceNorm backward)."""
import torch

from oracle import defectgan_ref as R
from oracle.bf16_emulation import bf16_storage, q


def _run(emulate):
    g_state = R.random_generator_state(ngf=16, hidden_nc=32, noise_weight=0.1)
    bg, labels, _ = R.synthetic_batch(2, 64, seed=0)
    gen = torch.Generator().manual_seed(7)
    noises = [torch.randn(s, generator=gen) for s in [(2, 1, 16, 16)] * 6 + [(2, 1, 32, 32), (2, 1, 64, 64)]]
    st = R.clone_state(g_state)
    for k in R.trainable_keys(st):
        st[k].requires_grad_(True)
    if emulate:
        with bf16_storage():
            out, prob = R.generator_forward(st, q(bg), labels, True, R.NoiseSource(noises))
            (out.square().mean() + prob.mean()).backward()
    else:
        out, prob = R.generator_forward(st, bg, labels, True, R.NoiseSource(noises))
        (out.square().mean() + prob.mean()).backward()
    return out.detach(), prob.detach(), {k: st[k].grad for k in R.trainable_keys(st) if st[k].grad is not None}


def test_bf16_storage_noise_level():
    o0, p0, g0 = _run(False)
    o1, p1, g1 = _run(True)
    rel = lambda a, b: ((a - b).abs().max() / b.abs().max()).item()  # noqa: E731
    assert rel(o1, o0) < 2e-2 and rel(p1, p0) < 2e-2
    a = torch.cat([g1[k].flatten() for k in g0])
    b = torch.cat([g0[k].flatten() for k in g0])
    rel_l2 = ((a - b).norm() / b.norm()).item()
    assert rel_l2 < 5e-2, rel_l2
    # ... yet element-wise max-norm bounds of a few per cent are NOT attainable for every parameter in BF16
    worst = max(rel(g1[k], g0[k]) for k in g0 if g0[k].norm() > 1e-3 * b.norm())
    assert worst > 5e-2, worst
    for k in g0:
        if g0[k].norm() > 1e-3 * b.norm():
            cos = torch.dot(g1[k].flatten(), g0[k].flatten()) / (g1[k].norm() * g0[k].norm())
            assert cos > 0.9, (k, cos.item())


def test_bf16_storage_noise_level_discriminator():
    """The discriminator (no norm layers, spectral-norm backward subtracts the rank-1 component): 4.7e-2 global."""
    def run(emulate):
        sd = R.clone_state(R.random_discriminator_state(ndf=8, image_size=128))
        xd, _, _ = R.synthetic_batch(2, 128, seed=1)
        for k in R.trainable_keys(sd):
            sd[k].requires_grad_(True)
        if emulate:
            with bf16_storage():
                s, c = R.discriminator_forward(sd, q(xd), True)
                (s.mean() + c.square().mean()).backward()
        else:
            s, c = R.discriminator_forward(sd, xd, True)
            (s.mean() + c.square().mean()).backward()
        return torch.cat([sd[k].grad.flatten() for k in R.trainable_keys(sd)])
    b, a = run(False), run(True)
    rel_l2 = ((a - b).norm() / b.norm()).item()
    assert 1e-2 < rel_l2 < 8e-2, rel_l2
