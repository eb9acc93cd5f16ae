"""Kernel-level parity through the C ABI: every CUDA entry point against a plain PyTorch FP32 CPU computation of the
same op on the same seeded inputs.  Inputs/weights are rounded to BF16 first so the comparison isolates the kernel
(FP32 accumulation, BF16 output rounding): tolerance 1e-2 of the output's max magnitude for BF16 outputs
(north-star: 2e-2 in BF16), 2e-3 for FP32 outputs."""
import json
import os

import pytest
import torch
import torch.nn.functional as F

from conftest import rel_err, ROOT

pytestmark = pytest.mark.gpu

TOL_BF16 = 1e-2
TOL_F32 = 2e-3


def _report(name, **kv):
    try:
        os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
        with open(os.path.join(ROOT, "gpurun_out", "kernel_parity.jsonl"), "a") as f:
            f.write(json.dumps(dict(test=name, **kv)) + "\n")
    except OSError:
        pass


def q(t):
    return t.bfloat16().float()


def to_dev(x_nchw):
    """fp32 NCHW CPU -> bf16 NHWC CUDA with the channel pitch padded to 8 (host-side, independent of our kernels)."""
    n, c, h, w = x_nchw.shape
    cp = (c + 7) // 8 * 8
    y = torch.zeros(n, h, w, cp)
    y[..., :c] = x_nchw.permute(0, 2, 3, 1)
    return y.bfloat16().cuda().contiguous()


def from_dev(y_nhwc, c):
    return y_nhwc.float().cpu()[..., :c].permute(0, 3, 1, 2).contiguous()


def ref_conv(x, w, b, stride, pad, pad_mode, size=None):
    if size is not None:
        x = F.interpolate(x, size=size, mode="nearest")
    if pad and pad_mode == "reflect":
        x = F.pad(x, (pad,) * 4, mode="reflect")
        pad = 0
    return F.conv2d(x, w, b, stride=stride, padding=pad)


CONV_CASES = {
    # name: (N, Cin, H, W, Cout, k, stride, pad, pad_mode, bias, resize_from)
    "res3x3_64": (2, 64, 16, 16, 64, 3, 1, 1, "reflect", False, None),
    "res3x3_256": (1, 256, 16, 16, 256, 3, 1, 1, "reflect", False, None),
    "enc4x4s2": (2, 64, 32, 32, 128, 4, 2, 1, "reflect", False, None),
    "stem7x7": (2, 3, 24, 24, 64, 7, 1, 3, "reflect", False, None),
    "spade_shared_1x1seg": (2, 6, 16, 16, 128, 3, 1, 1, "zeros", True, (1, 1)),
    "spade_shared_8x8seg": (2, 6, 16, 16, 128, 3, 1, 1, "zeros", True, (8, 8)),
    "spade_gb_512": (1, 128, 16, 16, 512, 3, 1, 1, "zeros", True, None),
    "heads_4": (2, 64, 16, 16, 4, 3, 1, 1, "reflect", False, None),
    "cls_valid_k2": (3, 256, 2, 2, 6, 2, 1, 0, "zeros", False, None),
    "ragged_m": (1, 64, 10, 10, 64, 3, 1, 1, "reflect", False, None),
    "d_conv0": (2, 3, 32, 32, 64, 4, 2, 1, "reflect", False, None),
    "d_src_1": (2, 128, 4, 4, 1, 3, 1, 1, "reflect", False, None),
    # TMA-im2col path (zero / no padding, Cin % 64 == 0, M % 128 == 0)
    "tma_valid3x3": (2, 64, 18, 18, 64, 3, 1, 0, "zeros", False, None),
    "tma_valid4x4s2": (2, 64, 34, 34, 128, 4, 2, 0, "zeros", False, None),
    "tma_zero3x3_256": (1, 256, 16, 16, 256, 3, 1, 1, "zeros", True, None),
    "tma_persistent_200tiles": (4, 64, 80, 80, 64, 3, 1, 1, "zeros", False, None),
    "tma_valid7x7": (2, 64, 22, 22, 64, 7, 1, 0, "zeros", False, None),
    "tma_dgrad_s2_to8": (2, 3, 34, 34, 64, 4, 2, 0, "zeros", False, None),   # D conv0 geometry: N tile mostly padding
    "tma_ragged_m100": (1, 64, 10, 10, 64, 3, 1, 1, "zeros", False, None),
    "tma_ragged_m75_s2": (3, 64, 12, 12, 72, 4, 2, 0, "zeros", False, None),
    "spade_cls_gb_5x5": (32, 128, 5, 5, 512, 3, 1, 1, "zeros", True, None),  # gamma|beta on the border-class map: M = 800
    "narrow_16_32": (2, 16, 12, 12, 32, 3, 1, 1, "reflect", False, None),   # reduced-width nets of the golden fixtures
    "narrow_s2_8_16": (2, 8, 16, 16, 16, 4, 2, 1, "reflect", False, None),
    # skinny kernels (conv_skinny.cu) at the geometries the model path produces (reflect padding pre-applied)
    "c8_stem_valid7x7": (2, 3, 70, 70, 64, 7, 1, 0, "zeros", False, None),
    "c8_zero3x3_nores": (2, 6, 16, 16, 128, 3, 1, 1, "zeros", True, None),
    "c8_spade_shared_big": (4, 6, 40, 40, 128, 3, 1, 1, "zeros", True, (1, 1)),
    "heads_valid_4": (2, 64, 18, 18, 4, 3, 1, 0, "zeros", False, None),
    "heads_valid_128_zero": (1, 128, 12, 12, 3, 3, 1, 1, "zeros", False, None),
    "heads_zero_64_wide": (3, 64, 9, 150, 4, 3, 1, 1, "zeros", False, None),   # 5 pixel chunks, zero frame, image restarts
    # operand-swapped wgrad (Cout <= 64 on Cin = 128 / 256, valid stride-1): x as the plain pixel matrix, dY im2col'ed per tap
    "wgrad_swap_128_64": (2, 128, 20, 24, 64, 3, 1, 0, "zeros", False, None),
    "wgrad_swap_256_60_ragged": (3, 256, 11, 13, 60, 3, 1, 0, "zeros", False, None),
    "wgrad_swap_128_64_k4": (2, 128, 12, 12, 64, 4, 1, 0, "zeros", False, None),
    "wgrad_swap_128_64_1x1": (2, 128, 9, 9, 64, 1, 1, 0, "zeros", False, None),
    "heads_valid_64_rows": (2, 64, 163, 34, 3, 3, 1, 0, "zeros", False, None),  # > 148 output rows per image: multi-row strips
    "d_cls_valid4x4": (4, 128, 4, 4, 6, 4, 1, 0, "zeros", False, None),
    # discriminator outputs at full width (conv_narrow_fwd_kernel: <= 8 output channels, K = taps * 2048)
    "d_cls_full_k32768": (4, 2048, 4, 4, 6, 4, 1, 0, "zeros", True, None),
    "d_src_full_valid3x3": (2, 2048, 6, 6, 1, 3, 1, 0, "zeros", False, None),
    "d_src_full_zero3x3": (2, 1024, 4, 4, 1, 3, 1, 1, "zeros", False, None),
}


def _make(case, seed=0):
    n, cin, h, w, cout, k, s, p, pm, bias, rs = CONV_CASES[case]
    g = torch.Generator().manual_seed(seed)
    hs, ws = rs if rs else (h, w)
    x = q(torch.randn(n, cin, hs, ws, generator=g))
    wt = q(torch.randn(cout, cin, k, k, generator=g) * (1.0 / (cin * k * k) ** 0.5))
    b = torch.randn(cout, generator=g) if bias else None
    return x, wt, b


@pytest.mark.parametrize("case", list(CONV_CASES))
def test_conv_fwd(case):
    from de_i2i_gan_b200 import ops
    n, cin, h, w, cout, k, s, p, pm, bias, rs = CONV_CASES[case]
    x, wt, b = _make(case)
    ref = ref_conv(x, wt, b, s, p, pm, (h, w) if rs else None)
    spec = ops.ConvSpec(cin, cout, k, s, p, pm)
    w_fwd, _ = ops.pack_weight(wt.cuda(), spec, need_dgrad=False)
    y = ops.conv2d_fwd(to_dev(x), w_fwd, spec, bias=b.cuda() if bias else None, logical_hw=(h, w) if rs else None)
    torch.cuda.synchronize()
    err = rel_err(from_dev(y, cout), ref)
    _report("conv_fwd/" + case, err=err)
    assert err < TOL_BF16, err
    if spec.cout_ld > cout and not bias:  # pad channels stay exactly zero
        assert y[..., cout:].abs().max().item() == 0.0


def test_reflect_pad_and_fold():
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(2)
    for pad in (1, 3):
        x = q(torch.randn(2, 16, 9, 11, generator=g)).requires_grad_(True)
        ref = F.pad(x, (pad,) * 4, mode="reflect")
        xp = ops.reflect_pad(to_dev(x.detach()), pad)
        assert torch.equal(from_dev(xp, 16), ref.detach())
        dyp = q(torch.randn(ref.shape, generator=g))
        (dx_ref,) = torch.autograd.grad(ref, x, dyp)
        dx = ops.reflect_fold(to_dev(dyp), pad)
        assert rel_err(from_dev(dx, 16), dx_ref) < TOL_BF16


def test_conv_tma_path_epilogue_matches_gather_path():
    """Same convolution through the TMA-im2col kernel and through the cp.async-gather kernel (tuning knob conv_path)."""
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(5)
    n, c, h, w = 2, 64, 16, 16
    x, wt = q(torch.randn(n, c, h, w, generator=g)), q(torch.randn(128, c, 3, 3, generator=g) / 24)
    noise = torch.randn(n, 1, h, w, generator=g).cuda().contiguous()
    nw, sc = torch.tensor([0.3]).cuda(), torch.tensor([0.7]).cuda()
    bias = torch.randn(128, generator=g).cuda()
    res = q(torch.randn(n, 128, h, w, generator=g))
    spec = ops.ConvSpec(c, 128, 3, 1, 1, "zeros")
    w_fwd, _ = ops.pack_weight(wt.cuda(), spec, need_dgrad=False)
    outs = []
    for path in (0, 1):
        ops.set_tuning("conv_path", path)
        try:
            outs.append(ops.conv2d_fwd(to_dev(x), w_fwd, spec, bias=bias, noise=noise, noise_weight=nw, residual=to_dev(res),
                                       scale=sc).float().cpu())
            outs.append(ops.conv2d_fwd(to_dev(x), w_fwd, spec, bias=bias, act="leaky_relu", out_fp32=True).cpu())
        finally:
            ops.set_tuning("conv_path", 0)
    ref = 0.7 * ref_conv(x, wt, None, 1, 1, "zeros") + bias.cpu()[None, :, None, None] + 0.3 * noise.cpu() + res
    assert rel_err(outs[0].permute(0, 3, 1, 2), ref) < TOL_BF16
    assert rel_err(outs[0], outs[2]) < 1e-2 and rel_err(outs[1], outs[3]) < 1e-3


def test_conv_fwd_fp32_out_and_epilogue():
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(5)
    n, c, h, w = 2, 64, 12, 12
    x, wt = q(torch.randn(n, c, h, w, generator=g)), q(torch.randn(c, c, 3, 3, generator=g) / 24)
    noise = torch.randn(n, 1, h, w, generator=g)
    nw = torch.tensor([0.3])
    res = q(torch.randn(n, c, h, w, generator=g))
    spec = ops.ConvSpec(c, c, 3, 1, 1, "reflect")
    w_fwd, _ = ops.pack_weight(wt.cuda(), spec, need_dgrad=False)
    conv = ref_conv(x, wt, None, 1, 1, "reflect")
    y = ops.conv2d_fwd(to_dev(x), w_fwd, spec, out_fp32=True)
    e1 = rel_err(y.cpu().permute(0, 3, 1, 2), conv)
    y2 = ops.conv2d_fwd(to_dev(x), w_fwd, spec, noise=noise.cuda().contiguous(), noise_weight=nw.cuda(), residual=to_dev(res),
                        act="leaky_relu")
    e2 = rel_err(from_dev(y2, c), F.leaky_relu(conv + 0.3 * noise, 0.2) + res)
    y3 = ops.conv2d_fwd(to_dev(x), w_fwd, spec, act="tanh", out_fp32=True)
    e3 = rel_err(y3.cpu().permute(0, 3, 1, 2), torch.tanh(conv))
    _report("conv_fwd/epilogue", fp32=e1, fused=e2, tanh=e3)
    assert e1 < TOL_F32 and e2 < TOL_BF16 and e3 < TOL_F32


def test_conv_heads_fwd_epilogue():
    """generator heads geometry (64 -> 4, 3x3 on a pre-padded map) through conv_heads_fwd_kernel: scale, bias, FP32 / BF16 out"""
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(6)
    n, c, h, w, co = 2, 64, 21, 37, 4
    x, wt = q(torch.randn(n, c, h, w, generator=g)), q(torch.randn(co, c, 3, 3, generator=g) / 24)
    b = torch.randn(co, generator=g)
    sc = torch.tensor([0.6])
    spec = ops.ConvSpec(c, co, 3, 1, 0, "zeros")
    w_fwd, _ = ops.pack_weight(wt.cuda(), spec, need_dgrad=False)
    conv = 0.6 * ref_conv(x, wt, None, 1, 0, "zeros") + b[None, :, None, None]
    y = ops.conv2d_fwd(to_dev(x), w_fwd, spec, bias=b.cuda(), scale=sc.cuda(), out_fp32=True)
    e1 = rel_err(y.cpu().permute(0, 3, 1, 2)[:, :co], conv)
    y2 = ops.conv2d_fwd(to_dev(x), w_fwd, spec, bias=b.cuda(), scale=sc.cuda(), act="tanh")
    e2 = rel_err(from_dev(y2, co), torch.tanh(conv))
    _report("conv_fwd/heads_epilogue", fp32=e1, tanh=e2)
    assert e1 < TOL_F32 and e2 < TOL_BF16, (e1, e2)


DGRAD_CASES = ["res3x3_64", "res3x3_256", "enc4x4s2", "spade_gb_512", "heads_4", "ragged_m", "d_src_1", "cls_valid_k2",
               "narrow_16_32", "narrow_s2_8_16", "tma_valid3x3", "tma_zero3x3_256", "tma_persistent_200tiles", "tma_valid4x4s2",
               "tma_valid7x7", "tma_ragged_m100", "tma_ragged_m75_s2", "tma_dgrad_s2_to8", "heads_valid_4",
               "heads_valid_128_zero", "d_cls_valid4x4", "c8_stem_valid7x7", "spade_cls_gb_5x5"]


@pytest.mark.parametrize("case", DGRAD_CASES)
def test_conv_dgrad(case):
    from de_i2i_gan_b200 import ops
    n, cin, h, w, cout, k, s, p, pm, bias, rs = CONV_CASES[case]
    x, wt, b = _make(case)
    x.requires_grad_(True)
    y = ref_conv(x, wt, None, s, p, pm)
    dy = q(torch.randn(y.shape, generator=torch.Generator().manual_seed(9)))
    (dx_ref,) = torch.autograd.grad(y, x, dy)
    spec = ops.ConvSpec(cin, cout, k, s, p, pm)
    _, w_dg = ops.pack_weight(wt.cuda(), spec)
    dx = ops.conv2d_dgrad(to_dev(dy), w_dg, spec, (h, w))
    torch.cuda.synchronize()
    err = rel_err(from_dev(dx, cin), dx_ref)
    _report("conv_dgrad/" + case, err=err)
    assert err < TOL_BF16, err


WGRAD_CASES = ["res3x3_64", "res3x3_256", "enc4x4s2", "stem7x7", "spade_shared_8x8seg", "spade_gb_512", "heads_4",
               "ragged_m", "d_conv0", "d_src_1", "cls_valid_k2", "narrow_16_32", "narrow_s2_8_16", "tma_valid3x3",
               "tma_zero3x3_256", "tma_persistent_200tiles", "tma_valid4x4s2", "tma_dgrad_s2_to8", "c8_stem_valid7x7",
               "c8_zero3x3_nores", "c8_spade_shared_big", "spade_shared_1x1seg", "heads_valid_4", "heads_valid_128_zero",
               "d_cls_valid4x4", "heads_zero_64_wide", "heads_valid_64_rows", "tma_ragged_m100", "tma_ragged_m75_s2",
               "spade_cls_gb_5x5", "wgrad_swap_128_64", "wgrad_swap_256_60_ragged", "wgrad_swap_128_64_k4", "wgrad_swap_128_64_1x1"]


@pytest.mark.parametrize("case", WGRAD_CASES)
def test_conv_wgrad(case):
    from de_i2i_gan_b200 import ops
    n, cin, h, w, cout, k, s, p, pm, bias, rs = CONV_CASES[case]
    x, wt, b = _make(case)
    wt.requires_grad_(True)
    y = ref_conv(x, wt, None, s, p, pm, (h, w) if rs else None)
    dy = q(torch.randn(y.shape, generator=torch.Generator().manual_seed(11)))
    (dw_ref,) = torch.autograd.grad(y, wt, dy)
    spec = ops.ConvSpec(cin, cout, k, s, p, pm)
    dw = torch.full(wt.shape, 7.0, device="cuda")
    ops.conv2d_wgrad(to_dev(x), to_dev(dy), dw, spec, logical_hw=(h, w) if rs else None)
    e1 = rel_err(dw.cpu(), dw_ref)
    ops.conv2d_wgrad(to_dev(x), to_dev(dy), dw, spec, accumulate=True, logical_hw=(h, w) if rs else None)
    e2 = rel_err(dw.cpu(), 2 * dw_ref)
    _report("conv_wgrad/" + case, err=e1, err_acc=e2, kernel=ops.last_kernel())
    assert e1 < TOL_F32 and e2 < TOL_F32, (e1, e2)
    if case.startswith("wgrad_swap"):
        assert "swap" in ops.last_kernel(), ops.last_kernel()
        ops.set_tuning("wgrad_swap", 0)  # and the ordinary operand roles on the same inputs
        try:
            dw2 = torch.zeros_like(dw)
            ops.conv2d_wgrad(to_dev(x), to_dev(dy), dw2, spec)
            assert "swap" not in ops.last_kernel()
        finally:
            ops.set_tuning("wgrad_swap", 1)
        assert rel_err(dw2.cpu(), dw_ref) < TOL_F32


def test_wgrad_spectral_backward_and_power_iteration():
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(3)
    cout, cin, k = 64, 64, 3
    w0 = torch.randn(cout, cin, k, k, generator=g) * 0.05
    u0 = F.normalize(torch.randn(cout, generator=g), dim=0)
    v0 = F.normalize(torch.randn(cin * k * k, generator=g), dim=0)
    # reference: one power iteration, then sigma (torch.nn.utils.spectral_norm, old API)
    wm = w0.reshape(cout, -1)
    v1 = F.normalize(wm.t() @ u0, dim=0, eps=1e-12)
    u1 = F.normalize(wm @ v1, dim=0, eps=1e-12)
    w0r = w0.clone().requires_grad_(True)
    sigma = torch.dot(u1, w0r.reshape(cout, -1) @ v1)
    wsn = w0r / sigma
    ud, vd = u0.cuda(), v0.cuda()
    inv = ops.spectral_norm(w0.cuda(), ud, vd, True)
    assert rel_err(ud.cpu(), u1) < 1e-4 and rel_err(vd.cpu(), v1) < 1e-4
    assert abs(inv[1].item() - sigma.item()) < 1e-4 * abs(sigma.item())
    inv_eval = ops.spectral_norm(w0.cuda(), ud, vd, False)
    assert abs(inv_eval[1].item() - sigma.item()) < 1e-4 * abs(sigma.item())
    # backward through W/sigma into W_orig
    x = q(torch.randn(2, cin, 12, 12, generator=g))
    y = ref_conv(x, wsn, None, 1, 1, "reflect")
    dy = q(torch.randn(y.shape, generator=g))
    (dw_ref,) = torch.autograd.grad(y, w0r, dy)
    spec = ops.ConvSpec(cin, cout, k, 1, 1, "reflect")
    dw = torch.zeros_like(w0, device="cuda")
    ops.conv2d_wgrad(to_dev(x), to_dev(dy), dw, spec, sn=(w0.cuda(), ud, vd, inv))
    err = rel_err(dw.cpu(), dw_ref)
    _report("wgrad_spectral", err=err)
    assert err < 5e-3, err
    # packed forward weight = W/sigma
    w_fwd, _ = ops.pack_weight(w0.cuda(), spec, inv_sigma=inv, need_dgrad=False)
    ref_pack = wsn.detach().permute(0, 2, 3, 1).reshape(cout, -1)
    assert rel_err(w_fwd.float().cpu()[:, : ref_pack.shape[1]], ref_pack) < 1e-2


def test_layout_roundtrip_and_fold():
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(1)
    x = torch.randn(2, 3, 9, 11, generator=g)
    d = ops.nchw_to_nhwc(x.cuda())
    assert d.shape == (2, 9, 11, 8) and d[..., 3:].abs().max().item() == 0
    assert rel_err(ops.nhwc_to_nchw(d, 3).cpu(), q(x)) == 0.0


@pytest.mark.parametrize("c,hw,per_sample", [(64, 16, False), (256, 8, True), (8, 32, False), (128, 20, True)])
def test_channel_stats(c, hw, per_sample):
    """(mean, biased variance) per channel / per (image, channel) - nn.BatchNorm2d / nn.InstanceNorm2d statistics."""
    from de_i2i_gan_b200 import ops
    x = q(torch.randn(3, c, hw, hw, generator=torch.Generator().manual_seed(2)) + 0.5)
    st = ops.channel_stats(to_dev(x), per_sample).cpu()
    dims = (2, 3) if per_sample else (0, 2, 3)
    assert rel_err(st[:, 0], x.mean(dims).reshape(st[:, 0].shape)) < 1e-5
    assert rel_err(st[:, 1], x.var(dims, unbiased=False).reshape(st[:, 1].shape)) < 1e-5


def test_channel_stats_large_mean_and_bitwise_repeatable():
    """|mean| >> std: sum(x^2)/n - mean^2 in FP32 would lose the variance (100^2 vs 0.25^2 leaves ~1 significant bit); the shifted /
    Chan-merged statistics keep it.  And the result is bit-identical run to run (no float atomics anywhere in the reduction)."""
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(5)
    x = q(100.0 + 0.25 * torch.randn(4, 64, 64, 64, generator=g))
    xd = to_dev(x)
    runs = [ops.channel_stats(xd, False).clone() for _ in range(3)] + [ops.channel_stats(xd, True).clone() for _ in range(2)]
    torch.cuda.synchronize()
    assert torch.equal(runs[0], runs[1]) and torch.equal(runs[0], runs[2]) and torch.equal(runs[3], runs[4])
    xr = x.double()
    var_ref = xr.var((0, 2, 3), unbiased=False)
    assert ((runs[0][0, 1].cpu().double() - var_ref).abs() / var_ref).max().item() < 1e-3
    assert ((runs[0][0, 0].cpu().double() - xr.mean((0, 2, 3))).abs()).max().item() < 1e-4


@pytest.mark.parametrize("kind,n,cin,h,cout,res", [("batch", 4, 64, 34, 64, False), ("sample", 3, 64, 18, 128, True),
                                                    ("batch", 2, 64, 21, 256, False), ("sample", 2, 128, 34, 64, False)])
def test_conv_epilogue_statistics(kind, n, cin, h, cout, res):
    """dei2i_conv_epilogue.stats_partial + dei2i_stats_finalize == dei2i_channel_stats of the stored output (what BatchNorm /
    InstanceNorm consume), incl. a ragged last tile (21x21 output) and the residual epilogue; bitwise repeatable."""
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(21)
    spec = ops.ConvSpec(cin, cout, 3, 1, 0, "zeros")
    x = to_dev(q(torch.randn(n, cin, h, h, generator=g)))
    wt = q(torch.randn(cout, cin, 3, 3, generator=g) * (1.0 / (cin * 9) ** 0.5))
    w_fwd, _ = ops.pack_weight(wt.cuda(), spec, need_dgrad=False)
    ho = h - 2
    r = to_dev(q(torch.randn(n, cout, ho, ho, generator=g))) if res else None
    assert ops.conv_fwd_is_tma(spec, n, h, h)
    y, st = ops.conv2d_fwd(x, w_fwd, spec, residual=r, want_stats=kind)
    y2, st2 = ops.conv2d_fwd(x, w_fwd, spec, residual=r, want_stats=kind)
    if kind == "sample" and (ho * ho) % 128 != 0:
        assert st is None
        return
    ref = ops.channel_stats(y, kind == "sample")
    torch.cuda.synchronize()
    assert torch.equal(st, st2) and torch.equal(y, y2)
    assert rel_err(st[:, 0], ref[:, 0]) < 1e-5 and rel_err(st[:, 1], ref[:, 1]) < 1e-5, (rel_err(st[:, 0], ref[:, 0]), rel_err(st[:, 1], ref[:, 1]))
    assert torch.equal(y, ops.conv2d_fwd(x, w_fwd, spec, residual=r))


@pytest.mark.parametrize("act,bias", [("leaky_relu", False), (None, True)])
def test_conv_epilogue_writes_consumer_padding(act, bias):
    """dei2i_conv_epilogue.out_pad: interior by the convolution + mirrored frame == reflect_pad of the plain output; and the
    activation backward with the fused fold == reflect_fold followed by the dense activation backward."""
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(22)
    n, cin, h, w, cout = 3, 64, 20, 14, 128
    spec = ops.ConvSpec(cin, cout, 4, 2, 0, "zeros")
    x = to_dev(q(torch.randn(n, cin, h, w, generator=g)))
    wt = q(torch.randn(cout, cin, 4, 4, generator=g) * (1.0 / (cin * 16) ** 0.5))
    b = torch.randn(cout, generator=g).cuda() if bias else None
    w_fwd, _ = ops.pack_weight(wt.cuda(), spec, need_dgrad=False)
    y0 = ops.conv2d_fwd(x, w_fwd, spec, bias=b, act=act)
    yp = ops.conv2d_fwd(x, w_fwd, spec, bias=b, act=act, out_pad=1)
    torch.cuda.synchronize()
    assert torch.equal(yp, ops.reflect_pad(y0, 1))
    dyp = to_dev(q(torch.randn(n, cout, y0.shape[1] + 2, y0.shape[2] + 2, generator=g)))
    dz_p, bg_p = ops.act_bwd(yp, dyp, act, want_dx=True, want_bias_grad=True, pad=1)
    dz_0, bg_0 = ops.act_bwd(y0, ops.reflect_fold(dyp, 1), act, want_dx=True, want_bias_grad=True)
    torch.cuda.synchronize()
    # the unfused path rounds the folded gradient to BF16 once more
    assert rel_err(dz_p.float(), dz_0.float()) < 1e-2 and rel_err(bg_p, bg_0) < 2e-3


@pytest.mark.parametrize("training,act,with_res", [(True, "leaky_relu", False), (True, None, True), (False, "leaky_relu", False)])
def test_batchnorm_fwd_bwd(training, act, with_res):
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(4)
    n, c, h = 3, 64, 12
    x = q(torch.randn(n, c, h, h, generator=g) * 1.5 + 0.3).requires_grad_(True)
    gamma = (1 + 0.1 * torch.randn(c, generator=g)).requires_grad_(True)
    beta = (0.1 * torch.randn(c, generator=g)).requires_grad_(True)
    rm, rv = torch.randn(c, generator=g) * 0.1, torch.rand(c, generator=g) + 0.5
    res = q(torch.randn(n, c, h, h, generator=g)) if with_res else None
    rm_ref, rv_ref = rm.clone(), rv.clone()
    y_ref = F.batch_norm(x, rm_ref, rv_ref, gamma, beta, training, 0.1, 1e-5)
    if act:
        y_ref = F.leaky_relu(y_ref, 0.2)
    if with_res:
        y_ref = y_ref + res
    dy = q(torch.randn(y_ref.shape, generator=g))
    dx_ref, dg_ref, db_ref = torch.autograd.grad(y_ref, (x, gamma, beta), dy)
    rmd, rvd = rm.cuda(), rv.cuda()
    xd = to_dev(x.detach())
    y, save = ops.bn_fwd(xd, gamma.detach().cuda(), beta.detach().cuda(), rmd, rvd, training, act=act,
                         residual=to_dev(res) if with_res else None)
    e_y = rel_err(from_dev(y, c), y_ref)
    assert e_y < TOL_BF16, e_y
    if training:
        assert rel_err(rmd.cpu(), rm_ref) < 1e-3 and rel_err(rvd.cpu(), rv_ref) < 1e-3
    dx, dgm, dbt = ops.bn_bwd(xd, to_dev(dy), gamma.detach().cuda(), beta.detach().cuda(), save, training, act=act)
    e = (rel_err(from_dev(dx, c), dx_ref), rel_err(dgm.cpu(), dg_ref), rel_err(dbt.cpu(), db_ref))
    _report(f"bn/{training}/{act}/{with_res}", y=e_y, dx=e[0], dgamma=e[1], dbeta=e[2])
    assert e[0] < TOL_BF16 and e[1] < 5e-3 and e[2] < 5e-3, e


def test_batchnorm_bwd_fused_fold_and_repeatable():
    """dei2i_bn_bwd with dy_pad == reflect_fold followed by the dense backward; dgamma / dbeta / dx bit-identical run to run."""
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(14)
    n, c, h, w = 3, 64, 11, 13
    x = to_dev(q(torch.randn(n, c, h, w, generator=g) * 1.5 + 0.3))
    gamma, beta = (1 + 0.3 * torch.randn(c, generator=g)).cuda(), (0.2 * torch.randn(c, generator=g)).cuda()
    rm, rv = torch.zeros(c, device="cuda"), torch.ones(c, device="cuda")
    _, save = ops.bn_fwd(x, gamma, beta, rm, rv, True, act="leaky_relu", out_pad=1)
    dyp = to_dev(q(torch.randn(n, c, h + 2, w + 2, generator=g)))
    a = ops.bn_bwd(x, dyp, gamma, beta, save, True, act="leaky_relu", dy_pad=1)
    b = ops.bn_bwd(x, dyp, gamma, beta, save, True, act="leaky_relu", dy_pad=1)
    ref = ops.bn_bwd(x, ops.reflect_fold(dyp, 1), gamma, beta, save, True, act="leaky_relu")
    torch.cuda.synchronize()
    for u, v in zip(a, b):
        assert torch.equal(u, v)
    assert rel_err(a[0].float(), ref[0].float()) < 1e-2 and rel_err(a[1], ref[1]) < 5e-3 and rel_err(a[2], ref[2]) < 5e-3
    # border folded in place (dei2i_reflect_fold_inplace), interior read through the padded pitch: the folded values are the ones
    # reflect_fold produces (same additions, rounded to BF16 once), so this matches the dense path bit for bit; pad = 2 as well
    for pad in (1, 2):
        dyp = to_dev(q(torch.randn(n, c, h + 2 * pad, w + 2 * pad, generator=g)))
        ref = ops.bn_bwd(x, ops.reflect_fold(dyp, pad), gamma, beta, save, True, act="leaky_relu")
        frame = dyp.clone()
        got = ops.bn_bwd(x, dyp, gamma, beta, save, True, act="leaky_relu", dy_pad=pad, fold_in_place=True)
        torch.cuda.synchronize()
        for u, v in zip(got, ref):
            assert torch.equal(u, v), pad
        inner = (slice(None), slice(pad, pad + h), slice(pad, pad + w))
        assert torch.equal(dyp[inner], ops.reflect_fold(frame, pad))  # the interior now holds the folded gradient ...
        mask = torch.ones_like(dyp, dtype=torch.bool)
        mask[inner] = False
        assert torch.equal(dyp[mask], frame[mask])                    # ... and the frame is untouched


@pytest.mark.parametrize("up", [1, 2])
def test_spade_fwd_bwd(up):
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(6)
    n, c, hx = 2, 64, 8
    h = hx * up
    x = q(torch.randn(n, c, hx, hx, generator=g) * 2 + 1).requires_grad_(True)
    gamma = q(torch.randn(n, c, h, h, generator=g) * 0.5).requires_grad_(True)
    beta = q(torch.randn(n, c, h, h, generator=g) * 0.5).requires_grad_(True)
    xu = F.interpolate(x, scale_factor=up, mode="nearest") if up > 1 else x
    out_ref = F.relu(F.instance_norm(xu, eps=1e-5) * (1 + gamma) + beta)
    dout = q(torch.randn(out_ref.shape, generator=g))
    dx_ref, dg_ref, db_ref = torch.autograd.grad(out_ref, (x, gamma, beta), dout)
    gb = torch.cat([gamma.detach(), beta.detach()], 1)
    xd, gbd = to_dev(x.detach()), to_dev(gb)
    out, mr = ops.spade_fwd(xd, gbd, up)
    e_out = rel_err(from_dev(out, c), out_ref)
    dgb, dx = ops.spade_bwd(xd, gbd, out, to_dev(dout), mr, up)
    dgb_h = from_dev(dgb, 2 * c)
    e = (rel_err(dx.float().cpu().permute(0, 3, 1, 2), dx_ref), rel_err(dgb_h[:, :c], dg_ref), rel_err(dgb_h[:, c:], db_ref))
    _report(f"spade/up{up}", out=e_out, dx=e[0], dgamma=e[1], dbeta=e[2])
    assert e_out < TOL_BF16 and max(e) < 2e-2, (e_out, e)


@pytest.mark.parametrize("use_sn,n,h,w", [(True, 2, 24, 32), (False, 3, 16, 16)])
def test_rowpack_stem_conv(use_sn, n, h, w):
    """Generator stem (3->64, 7x7, 'same' reflect, spectral norm) through the row-packed 7x1 TMA convolution vs nn.Conv2d."""
    import copy
    from de_i2i_gan_b200 import functional as Fn, ops
    torch.manual_seed(3)
    conv = torch.nn.Conv2d(3, 64, 7, padding="same", padding_mode="reflect", bias=False)
    with torch.no_grad():
        conv.weight.copy_(q(conv.weight))
    if use_sn:
        conv = torch.nn.utils.spectral_norm(conv)
    dev = copy.deepcopy(conv).cuda()
    g = torch.Generator().manual_seed(4)
    x = q(torch.randn(n, 3, h, w, generator=g)).requires_grad_(True)
    y_ref = conv(x)
    dy = q(torch.randn(y_ref.shape, generator=g))
    wref = conv.weight_orig if use_sn else conv.weight
    dx_ref, dw_ref = torch.autograd.grad(y_ref, (x, wref), dy)
    spec = ops.ConvSpec(3, 64, 7, 1, 3, "reflect")
    assert Fn.rowpack_applicable(spec)
    xd = x.detach().cuda().requires_grad_(True)
    y = Fn.rowpack_conv(xd, dev, spec, True)
    wdev = dev.weight_orig if use_sn else dev.weight
    dxd, dwd = torch.autograd.grad(y.t, (xd, wdev), to_dev(dy))
    torch.cuda.synchronize()
    e = (rel_err(from_dev(y.t, 64), y_ref), rel_err(dxd, dx_ref), rel_err(dwd, dw_ref))
    _report(f"rowpack_stem/sn{int(use_sn)}", y=e[0], dx=e[1], dw=e[2])
    assert e[0] < TOL_BF16 and e[1] < TOL_BF16 and e[2] < TOL_BF16, e
    if use_sn:
        assert rel_err(dev.weight_u, conv.weight_u) < 1e-4 and rel_err(dev.weight_v, conv.weight_v) < 1e-4


def test_rowpack_roundtrip_adjoint():
    """rowpack_bwd is the adjoint of rowpack_fwd: <P x, z> == <x, P^T z>."""
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(8)
    x = q(torch.randn(2, 3, 9, 11, generator=g))
    xp = ops.rowpack_fwd(x.cuda(), 7, 3).float()
    ref = F.pad(x, (3, 3, 3, 3), mode="reflect")
    for s_ in range(7):
        got = xp[:, :, :, s_ * 8: s_ * 8 + 3].permute(0, 3, 1, 2).cpu()
        assert torch.equal(got, ref[:, :, :, s_: s_ + 11])
    assert xp[..., 56:].abs().max().item() == 0 and xp[..., 3:8].abs().max().item() == 0
    z = q(torch.randn(xp.shape, generator=g))
    pt = ops.rowpack_bwd(z.to(torch.bfloat16).cuda().contiguous(), 3, 7, 3)
    zz = z.clone()
    zz[..., 56:] = 0
    zz.view(2, 15, 11, 8, 8)[..., 3:] = 0
    lhs = (xp.cpu() * zz).sum().item()
    rhs = (x * pt.cpu()).sum().item()
    assert abs(lhs - rhs) < 1e-2 * max(1.0, abs(lhs)), (lhs, rhs)


def _cls_index(n):
    return torch.tensor([i if i < 2 else (i - n + 5 if i >= n - 2 else 2) for i in range(n)])


@pytest.mark.parametrize("up,hx,wx,c", [(1, 8, 8, 64), (2, 8, 6, 64), (1, 5, 7, 256), (2, 16, 16, 128)])
def test_spade_cls_fwd_bwd(up, hx, wx, c):
    """SPADE over the 5x5 border-class map vs plain torch on the gathered full-resolution gamma / beta."""
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(16)
    n = 3
    h, w = hx * up, wx * up
    x = q(torch.randn(n, c, hx, wx, generator=g) * 2 + 1).requires_grad_(True)
    gamma5 = q(torch.randn(n, c, 5, 5, generator=g) * 0.5).requires_grad_(True)
    beta5 = q(torch.randn(n, c, 5, 5, generator=g) * 0.5).requires_grad_(True)
    ih, iw = _cls_index(h), _cls_index(w)
    gamma = gamma5[:, :, ih][:, :, :, iw]
    beta = beta5[:, :, ih][:, :, :, iw]
    xu = F.interpolate(x, scale_factor=up, mode="nearest") if up > 1 else x
    out_ref = F.relu(F.instance_norm(xu, eps=1e-5) * (1 + gamma) + beta)
    dout = q(torch.randn(out_ref.shape, generator=g))
    dx_ref, dg_ref, db_ref = torch.autograd.grad(out_ref, (x, gamma5, beta5), dout)
    xd, gbd = to_dev(x.detach()), to_dev(torch.cat([gamma5.detach(), beta5.detach()], 1))
    out, mr = ops.spade_cls_fwd(xd, gbd, up)
    e_out = rel_err(from_dev(out, c), out_ref)
    dgb5, dx = ops.spade_cls_bwd(xd, gbd, out, to_dev(dout), mr, up)
    torch.cuda.synchronize()
    dgb_h = dgb5.cpu().permute(0, 3, 1, 2)
    e = (rel_err(dx.float().cpu().permute(0, 3, 1, 2), dx_ref), rel_err(dgb_h[:, :c], dg_ref), rel_err(dgb_h[:, c:], db_ref))
    _report(f"spade_cls/up{up}/{hx}x{wx}x{c}", out=e_out, dx=e[0], dgamma=e[1], dbeta=e[2])
    assert e_out < TOL_BF16 and e[0] < 2e-2 and max(e[1:]) < 5e-3, (e_out, e)


@pytest.mark.parametrize("up,pad", [(1, 1), (2, 1)])
def test_spade_cls_fused_reflect_pad(up, pad):
    """out_pad: the class-map SPADE writes its consumer's reflect padding and folds the padded gradient - equal to the
    unfused kernels followed by reflect_pad / preceded by reflect_fold."""
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(17)
    n, c, hx, wx = 2, 64, 7, 9
    h, w = hx * up, wx * up
    xd = to_dev(q(torch.randn(n, c, hx, wx, generator=g) * 2 + 1))
    gbd = to_dev(q(torch.randn(n, 2 * c, 5, 5, generator=g) * 0.5))
    out0, mr = ops.spade_cls_fwd(xd, gbd, up)
    outp, mrp = ops.spade_cls_fwd(xd, gbd, up, out_pad=pad)
    assert torch.equal(outp, ops.reflect_pad(out0, pad)) and torch.equal(mr, mrp)
    doutp = to_dev(q(torch.randn(n, c, h + 2 * pad, w + 2 * pad, generator=g)))
    dgb_p, dx_p = ops.spade_cls_bwd(xd, gbd, outp, doutp, mr, up, out_pad=pad)
    dgb_0, dx_0 = ops.spade_cls_bwd(xd, gbd, out0, ops.reflect_fold(doutp, pad), mr, up)
    torch.cuda.synchronize()
    # the unfused path rounds the folded gradient to BF16 once more
    assert rel_err(dx_p.float(), dx_0.float()) < 1e-2 and rel_err(dgb_p, dgb_0) < 1e-2
    dgb_q, dx_q = ops.spade_cls_bwd(xd, gbd, outp, doutp, mr, up, out_pad=pad)
    torch.cuda.synchronize()
    assert torch.equal(dgb_p, dgb_q) and torch.equal(dx_p, dx_q)  # deterministic: no atomics in the class sums


def test_spade_module_class_map_equals_dense():
    """SPADE module on a 1x1 label map: border-class path == dense path == the reference formula (normalization.py:24-37)."""
    from de_i2i_gan_b200.models.networks import normalization as nz
    g = torch.Generator().manual_seed(21)
    torch.manual_seed(5)
    n, c, hx, up = 2, 64, 12, 2
    mod = nz.SPADE(6, c, hidden_nc=128, kernel_size=(3, 3), padding="same", norm_layer=torch.nn.InstanceNorm2d).cuda()
    with torch.no_grad():
        for p_ in mod.parameters():
            p_.copy_(q(p_.detach().cpu() * 3).cuda())
    x = q(torch.randn(n, c, hx, hx, generator=g))
    lab = (torch.rand(n, 6, 1, 1, generator=g) > 0.5).float()
    dout = q(torch.randn(n, c, hx * up, hx * up, generator=g))

    def run(flag):
        nz.CONSTANT_LABEL_CLASSES = flag
        xg = x.cuda().requires_grad_(True)
        mod.zero_grad()
        out = mod(xg, lab.cuda(), up=up, relu=True)
        out.backward(dout.cuda())
        return out.detach().cpu(), xg.grad.cpu(), [p_.grad.detach().cpu().clone() for p_ in mod.parameters()]

    try:
        o1, dx1, gp1 = run(True)
        o0, dx0, gp0 = run(False)
    finally:
        nz.CONSTANT_LABEL_CLASSES = True
    # reference formula in fp32 on the CPU
    xr = x.clone().requires_grad_(True)
    ps = [p_.detach().cpu().clone().requires_grad_(True) for p_ in mod.parameters()]
    names = [k for k, _ in mod.named_parameters()]
    pd = dict(zip(names, ps))
    xu = F.interpolate(xr, scale_factor=up, mode="nearest")
    seg = F.interpolate(lab, size=xu.shape[2:], mode="nearest")
    ste = lambda t: t + (q(t.detach()) - t.detach())  # BF16 storage of the intermediates (straight-through): keeps ReLU masks equal
    actv = ste(F.relu(F.conv2d(seg, pd["mlp_shared.0.weight"], pd["mlp_shared.0.bias"], padding=1)))
    gamma = ste(F.conv2d(actv, pd["mlp_gamma.weight"], pd["mlp_gamma.bias"], padding=1))
    beta = ste(F.conv2d(actv, pd["mlp_beta.weight"], pd["mlp_beta.bias"], padding=1))
    out_ref = F.relu(F.instance_norm(xu, eps=1e-5) * (1 + gamma) + beta)
    grads = torch.autograd.grad(out_ref, [xr] + ps, dout)
    l2 = lambda a, b: ((a - b).norm() / b.norm()).item()  # a flipped ReLU mask is a one-pixel 100 % error: L2, not max
    e_cls = [rel_err(o1, out_ref), l2(dx1, grads[0])] + [l2(a, b) for a, b in zip(gp1, grads[1:])]
    e_dense = [rel_err(o0, out_ref), l2(dx0, grads[0])] + [l2(a, b) for a, b in zip(gp0, grads[1:])]
    _report("spade_module/class_map_vs_ref", errs=e_cls)
    _report("spade_module/dense_vs_ref", errs=e_dense)
    print("cls", e_cls, "dense", e_dense, "cls-vs-dense dx", rel_err(dx1, dx0))
    assert max(e_cls) < 2e-2, e_cls
    assert max(e_dense) < 2e-2, e_dense
    assert rel_err(o1, o0) < TOL_BF16


@pytest.mark.parametrize("act", ["leaky_relu", "relu", "tanh", "sigmoid", None])
def test_act_bwd_and_bias_grad(act):
    from de_i2i_gan_b200 import ops
    from oracle.defectgan_ref import act as ref_act
    g = torch.Generator().manual_seed(7)
    z = torch.randn(2, 64, 9, 9, generator=g).requires_grad_(True)
    y = ref_act(z, act)
    dy = q(torch.randn(y.shape, generator=g))
    (dz_ref,) = torch.autograd.grad(y, z, dy)
    yq = to_dev(y.detach())
    dx, bg = ops.act_bwd(yq, to_dev(dy), act, True, True)
    e1, e2 = rel_err(from_dev(dx, 64), dz_ref), rel_err(bg.cpu(), dz_ref.sum((0, 2, 3)))
    _report(f"act_bwd/{act}", dx=e1, bias=e2)
    assert e1 < 2e-2 and e2 < 2e-2


def test_noise_wgrad():
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(8)
    dy = q(torch.randn(2, 64, 10, 10, generator=g))
    noise = torch.randn(2, 1, 10, 10, generator=g)
    ref = (dy * noise).sum()
    dw = torch.zeros(1, device="cuda")
    ops.noise_wgrad(to_dev(dy), noise.cuda().contiguous(), dw, 64)
    ops.noise_wgrad(to_dev(dy), noise.cuda().contiguous(), dw, 64, accumulate=True)
    assert abs(dw.item() - 2 * ref.item()) < 1e-3 * abs(ref.item()) + 1e-3


@pytest.mark.parametrize("cycle", [False, True])
def test_compose_fwd_bwd(cycle):
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(10)
    n, h = 2, 12
    hd = torch.randn(n, h, h, 8, generator=g)
    x = torch.randn(n, 3, h, h, generator=g).clamp(-1, 1).requires_grad_(True)
    hdr = hd.clone().requires_grad_(True)
    fg = torch.tanh(hdr[..., :3]).permute(0, 3, 1, 2)
    p = torch.sigmoid(hdr[..., 3:4]).permute(0, 3, 1, 2)
    out_ref = fg if cycle else x * (1 - p) + fg * p
    dout, dprob = torch.randn(n, 3, h, h, generator=g), torch.randn(n, 1, h, h, generator=g)
    grads = torch.autograd.grad((out_ref, p), (hdr, x), (dout, dprob), allow_unused=True)
    out, prob, onhwc = ops.compose_fwd(hd.cuda(), x.detach().cuda(), cycle)
    assert rel_err(out.cpu(), out_ref) < 1e-4 and rel_err(prob.cpu(), p) < 1e-4
    assert rel_err(onhwc.float().cpu()[..., :3].permute(0, 3, 1, 2), out_ref) < 1e-2
    dhd, dx = ops.compose_bwd(hd.cuda(), x.detach().cuda(), dout.cuda(), dprob.cuda(), cycle)
    assert rel_err(dhd.float().cpu()[..., :4], grads[0][..., :4]) < 1e-2
    if not cycle:
        assert rel_err(dx.cpu(), grads[1]) < 1e-4


def test_losses_and_adam():
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(12)
    z = (torch.randn(4, 1, 4, 4, generator=g) * 3).requires_grad_(True)
    for tc in (0.0, 1.0):
        ref = F.binary_cross_entropy_with_logits(z, torch.full_like(z, tc))
        (gr,) = torch.autograd.grad(ref, z)
        loss, grad = ops.bce_logits(z.detach().cuda(), target_const=tc, grad_scale=0.5)
        assert abs(loss.item() - ref.item()) < 1e-5 and rel_err(grad.cpu(), 0.5 * gr) < 1e-4
    t = (torch.rand(4, 6, generator=g) > 0.5).float()
    zz = torch.randn(4, 6, generator=g).requires_grad_(True)
    ref = F.binary_cross_entropy_with_logits(zz, t)
    (gr,) = torch.autograd.grad(ref, zz)
    loss, grad = ops.bce_logits(zz.detach().cuda(), targets=t.cuda())
    assert abs(loss.item() - ref.item()) < 1e-5 and rel_err(grad.cpu(), gr) < 1e-4
    a = torch.randn(2, 3, 40, 40, generator=g).requires_grad_(True)
    b = torch.randn(2, 3, 40, 40, generator=g).requires_grad_(True)
    ref = F.l1_loss(a, b)
    ga_ref, gb_ref = torch.autograd.grad(ref, (a, b))
    ga, gb = torch.zeros_like(a, device="cuda"), torch.zeros_like(b, device="cuda")
    loss = ops.l1(a.detach().cuda(), b.detach().cuda(), ga, gb, grad_scale=2.0)
    assert abs(loss.item() - ref.item()) < 1e-5
    assert rel_err(ga.cpu(), 2 * ga_ref) < 1e-5 and rel_err(gb.cpu(), 2 * gb_ref) < 1e-5
    loss0 = ops.l1(a.detach().cuda())
    assert abs(loss0.item() - a.abs().mean().item()) < 1e-5
    # Adam vs torch.optim.Adam over 3 steps
    p = torch.randn(1000, generator=g)
    pt = p.clone().requires_grad_(True)
    opt = torch.optim.Adam([pt], lr=2e-4, betas=(0.5, 0.999))
    pd, m, v = p.cuda(), torch.zeros(1000, device="cuda"), torch.zeros(1000, device="cuda")
    for step in range(1, 4):
        gr = torch.randn(1000, generator=g)
        pt.grad = gr.clone()
        opt.step()
        ops.adam(pd, gr.cuda(), m, v, 2e-4, 0.5, 0.999, 1e-8, step)
    assert (pd.cpu() - pt.detach()).abs().max().item() < 1e-6


def test_errors_are_reported_not_fatal():
    from de_i2i_gan_b200 import ops
    spec = ops.ConvSpec(64, 64, 3, 1, 1, "reflect")
    x = torch.zeros(1, 1, 1, 64, dtype=torch.bfloat16, device="cuda")  # reflect pad 1 needs H,W > 1
    w = torch.zeros(64, spec.kf, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(RuntimeError, match="reflect"):
        ops.conv2d_fwd(x, w, spec)


@pytest.mark.parametrize("case", ["tma_valid3x3", "tma_valid4x4s2", "tma_ragged_m100", "tma_ragged_m75_s2", "tma_persistent_200tiles"])
def test_conv_tma_two_subtiles(case):
    """conv_tma_kernel with MS = 2 (two 128-pixel A tiles per B tile; picked automatically only for large M): forward and dgrad."""
    from de_i2i_gan_b200 import ops
    n, cin, h, w, cout, k, s, p, pm, bias, rs = CONV_CASES[case]
    x, wt, b = _make(case)
    x.requires_grad_(True)
    ref = ref_conv(x, wt, b, s, p, pm)
    dy = q(torch.randn(ref.shape, generator=torch.Generator().manual_seed(9)))
    (dx_ref,) = torch.autograd.grad(ref, x, dy)
    spec = ops.ConvSpec(cin, cout, k, s, p, pm)
    w_fwd, w_dg = ops.pack_weight(wt.cuda(), spec)
    ops.set_tuning("tma_ms", 2)
    try:
        y = ops.conv2d_fwd(to_dev(x.detach()), w_fwd, spec, bias=b.cuda() if bias else None)
        dx = ops.conv2d_dgrad(to_dev(dy), w_dg, spec, (h, w))
        torch.cuda.synchronize()
    finally:
        ops.set_tuning("tma_ms", 0)
    e1, e2 = rel_err(from_dev(y, cout), ref.detach()), rel_err(from_dev(dx, cin), dx_ref)
    _report("conv_tma_ms2/" + case, err_fwd=e1, err_dgrad=e2)
    assert e1 < TOL_BF16 and e2 < TOL_BF16, (e1, e2)


@pytest.mark.parametrize("training", [True, False])
def test_spectral_norm_multi_matches_single(training):
    """dei2i_spectral_norm_multi (four launches for a whole network) is bit-identical to the per-weight entry point."""
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(11)
    shapes = [(64, 3, 7, 7), (128, 64, 4, 4), (256, 256, 3, 3), (2304, 40, 1, 1), (8, 2048, 3, 3)] + [(16 + i, 8 + i, 3, 3) for i in range(45)]
    ws = [torch.randn(s_, generator=g).cuda() for s_ in shapes]
    us = [F.normalize(torch.randn(s_[0], generator=g), dim=0).cuda() for s_ in shapes]
    vs = [F.normalize(torch.randn(s_[1] * s_[2] * s_[3], generator=g), dim=0).cuda() for s_ in shapes]
    u1, v1 = [u.clone() for u in us], [v.clone() for v in vs]
    ref = [ops.spectral_norm(w, u, v, training).clone() for w, u, v in zip(ws, u1, v1)]
    u2, v2 = [u.clone() for u in us], [v.clone() for v in vs]
    got = ops.spectral_norm_multi(list(zip(ws, u2, v2)), training)
    torch.cuda.synchronize()
    for i, (inv, su, sv) in enumerate(got):
        assert torch.equal(inv, ref[i]), (i, inv, ref[i])
        assert torch.equal(u2[i], u1[i]) and torch.equal(v2[i], v1[i]), i
        assert torch.equal(su, u2[i]) and torch.equal(sv, v2[i]), i
        if not training:
            assert torch.equal(u2[i], us[i]) and torch.equal(v2[i], vs[i])
    # sigma is the top singular value direction estimate: sanity against torch for the first weight
    w0 = ws[2].reshape(256, -1)
    assert abs(got[2][0][1].item() - torch.dot(u2[2], w0 @ v2[2]).item()) < 1e-3 * got[2][0][1].item()


def test_spectral_norm_prefetch_in_networks():
    """The generator / discriminator forward consumes every prefetched entry and leaves u, v exactly as the per-layer path."""
    from pathlib import Path
    import copy
    from de_i2i_gan_b200 import functional as Fn, options
    from de_i2i_gan_b200.models.networks.generator import DefectGanGenerator
    from de_i2i_gan_b200.models.networks.discriminator import DefectGanDiscriminator
    opt = options.readme_recipe(image_size=128, batch_size=2, ckpt_dir=Path("/tmp/dei2i_sn_test"), ngf=8, ndf=8)
    torch.manual_seed(5)
    G0, D0 = DefectGanGenerator(opt).cuda(), DefectGanDiscriminator(opt).cuda()
    x = torch.randn(2, 3, 128, 128, device="cuda").clamp_(-1, 1)
    lab = torch.zeros(2, opt.label_nc, device="cuda")
    lab[:, 1] = 1
    state = {}
    for multi in (True, False):
        Fn.SN_MULTI = multi
        try:
            G, D = copy.deepcopy(G0), copy.deepcopy(D0)
            for _ in range(3):  # the first forward records which wrapped convolutions run, the later ones prefetch them
                with Fn.weight_cache():
                    out = G(x, lab)
                    D(x)
            assert not Fn._sn_ready, "prefetched spectral-norm entries left unconsumed"
            assert not multi or (len(G._sn_used) == 17 and len(D._sn_used) == 6)
        finally:
            Fn.SN_MULTI = True
        state[multi] = {f"{i}.{n}": b.clone() for i, net in enumerate((G, D)) for n, b in net.named_buffers()
                        if n.endswith("weight_u") or n.endswith("weight_v")}
    assert len(state[True]) == 2 * (20 + 6)
    for n, b in state[True].items():
        assert torch.equal(b, state[False][n]), n
        if "conv_s" in n:  # learned shortcuts that never run keep their initial vectors
            assert torch.equal(b, dict(G0.named_buffers())[n[2:]].to(b.device))


@pytest.mark.parametrize("pad,act,with_res", [(1, "leaky_relu", False), (3, "relu", True), (1, None, True)])
def test_bn_fwd_writes_consumer_padding(pad, act, with_res):
    """dei2i_bn_fwd_pad == reflect_pad(dei2i_bn_fwd): output, saved statistics and running statistics."""
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(12)
    n, c, h, w = 3, 64, 9, 11
    x = to_dev(q(torch.randn(n, c, h, w, generator=g) * 1.5 + 0.3))
    res = to_dev(q(torch.randn(n, c, h, w, generator=g))) if with_res else None
    gamma, beta = (1 + 0.3 * torch.randn(c, generator=g)).cuda(), (0.2 * torch.randn(c, generator=g)).cuda()
    rm0, rv0 = torch.zeros(c, device="cuda"), torch.ones(c, device="cuda")
    rm1, rv1 = rm0.clone(), rv0.clone()
    y0, save0 = ops.bn_fwd(x, gamma, beta, rm0, rv0, True, act=act, residual=res)
    yp, savep = ops.bn_fwd(x, gamma, beta, rm1, rv1, True, act=act, residual=res, out_pad=pad)
    torch.cuda.synchronize()
    ref = ops.reflect_pad(y0, pad)
    assert yp.shape == ref.shape
    assert rel_err(yp.float(), ref.float()) < 1e-2  # one BF16 ulp where the two kernels contract the FMA differently
    assert (yp != ref).float().mean().item() < 0.01
    assert torch.allclose(savep, save0, rtol=1e-6, atol=1e-7) and torch.allclose(rm1, rm0, rtol=1e-6, atol=1e-7)
    assert torch.allclose(rv1, rv0, rtol=1e-6, atol=1e-7)


def test_pack_weight_multi_matches_per_weight_packing():
    """dei2i_pack_weight_multi (one launch per 32 weights, incl. the row-wise concatenation of two weights) against dei2i_pack_weight:
    bit-identical forward and dgrad matrices; 40 items -> two launches."""
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(21)
    entries, want = [], []
    shapes = [(64, 3, 7, 7), (128, 64, 4, 4), (256, 256, 3, 3), (6, 128, 3, 3), (1, 64, 3, 3), (48, 24, 1, 1), (64, 64, 7, 1)] * 5
    for i, (co, ci, kh, kw) in enumerate(shapes):
        w = torch.randn(co, ci, kh, kw, generator=g).cuda()
        spec = ops.ConvSpec(ci, co, kh, 1, 0, "zeros", kw if kw != kh else 0)
        entries.append((w, None, spec))
        want.append(ops.pack_weight(w, spec))
    for ca, cb, ci in ((256, 256, 128), (3, 1, 64), (5, 120, 16), (64, 64, 128), (8, 8, 8)):
        wa, wb = torch.randn(ca, ci, 3, 3, generator=g).cuda(), torch.randn(cb, ci, 3, 3, generator=g).cuda()
        spec = ops.ConvSpec(ci, ca + cb, 3, 1, 1, "zeros")
        entries.append((wa, wb, spec))
        want.append(ops.pack_weight(torch.cat([wa, wb], 0).contiguous(), spec))
    l0 = ops.launch_count()
    got = ops.pack_weight_multi(entries)
    assert ops.launch_count() - l0 == 2
    torch.cuda.synchronize()
    for (gf, gd), (wf, wd), (w, _, spec) in zip(got, want, entries):
        assert gf.shape == wf.shape and gd.shape == wd.shape
        assert torch.equal(gf, wf) and torch.equal(gd, wd), spec
        assert gf.data_ptr() % 256 == 0 and gd.data_ptr() % 256 == 0
    # batches without a > 16-tap filter take the shared-memory-tiled kernel: same bits
    small = [(e, w_) for e, w_ in zip(entries, want) if e[2].kh * e[2].kw_ <= 16]
    for (spec_extra) in ((3, 64, 4), (6, 128, 3), (100, 20, 2), (8, 1, 4)):  # ragged channel counts on the tiled path
        ci, co, k = spec_extra
        w = torch.randn(co, ci, k, k, generator=g).cuda()
        sp = ops.ConvSpec(ci, co, k, 1, 0, "zeros")
        small.append(((w, None, sp), ops.pack_weight(w, sp)))
    got = ops.pack_weight_multi([e for e, _ in small])
    torch.cuda.synchronize()
    for (gf, gd), (e, (wf, wd)) in zip(got, small):
        assert torch.equal(gf, wf) and torch.equal(gd, wd), e[2]


@pytest.mark.parametrize("n,ho,w,kh", [(4, 40, 256, 7), (2, 20, 128, 7), (3, 9, 256, 3), (1, 150, 128, 7), (33, 7, 128, 7)])
def test_conv_ring_khx1_forward_and_dgrad(n, ho, w, kh):
    """conv_ring_kernel (kh x 1 filters over 64 channels, the row-packed stem): forward and dgrad against torch and against the generic
    TMA kernel on the same inputs.  Shapes make CTAs own one row, several rows, and ranges that cross image / column-block borders."""
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(100 + n + ho)
    h = ho + kh - 1
    x = q(torch.randn(n, 64, h, w, generator=g))
    wt = q(torch.randn(64, 64, kh, 1, generator=g) / (8 * kh ** 0.5))
    sc = torch.tensor([0.7], device="cuda")
    spec = ops.ConvSpec(64, 64, kh, 1, 0, "zeros", 1)
    w_fwd, w_dg = ops.pack_weight(wt.cuda(), spec)
    xr = x.clone().requires_grad_(True)
    yr = 0.7 * F.conv2d(xr, wt)
    dy = q(torch.randn(yr.shape, generator=g))
    (dx_ref,) = torch.autograd.grad(yr, xr, dy)
    outs = {}
    for ring in (1, 0):
        ops.set_tuning("conv_ring", ring)
        try:
            y = ops.conv2d_fwd(to_dev(x), w_fwd, spec, scale=sc)
            k_f = ops.last_kernel()
            dx = ops.conv2d_dgrad(to_dev(dy), w_dg, spec, (h, w), scale=sc)
            k_d = ops.last_kernel()
        finally:
            ops.set_tuning("conv_ring", 1)
        torch.cuda.synchronize()
        assert ("conv_ring" in k_f) == bool(ring) and ("conv_ring" in k_d) == bool(ring), (k_f, k_d)
        outs[ring] = (from_dev(y, 64), from_dev(dx, 64))
    # weight gradient on the row-ring wgrad kernel (mirrored ring, one accumulator per CTA) vs torch and vs the generic kernel
    wr = wt.clone().requires_grad_(True)
    (dw_ref,) = torch.autograd.grad(F.conv2d(x, wr), wr, dy)
    dws = {}
    for ring in (1, 0):
        ops.set_tuning("wgrad_ring", ring)
        try:
            dw = torch.full(wt.shape, 3.0, device="cuda")
            ops.conv2d_wgrad(to_dev(x), to_dev(dy), dw, spec)
            k_w = ops.last_kernel()
            ops.conv2d_wgrad(to_dev(x), to_dev(dy), dw, spec, accumulate=True)
        finally:
            ops.set_tuning("wgrad_ring", 1)
        torch.cuda.synchronize()
        assert ("wgrad_ring" in k_w) == bool(ring), k_w
        dws[ring] = dw.cpu()
    e_w = rel_err(dws[1], 2 * dw_ref)
    assert e_w < TOL_F32 and rel_err(dws[1], dws[0]) < TOL_F32, (e_w, rel_err(dws[1], dws[0]))
    e_f, e_d = rel_err(outs[1][0], yr.detach()), rel_err(outs[1][1], dx_ref)
    _report(f"conv_ring/{n}x{ho}x{w}k{kh}", fwd=e_f, dgrad=e_d, wgrad=e_w, vs_generic_fwd=rel_err(outs[1][0], outs[0][0]), vs_generic_dgrad=rel_err(outs[1][1], outs[0][1]))
    assert e_f < TOL_BF16 and e_d < TOL_BF16, (e_f, e_d)
    assert torch.equal(outs[1][0], outs[0][0]) or rel_err(outs[1][0], outs[0][0]) < 4e-3
    assert torch.equal(outs[1][1], outs[0][1]) or rel_err(outs[1][1], outs[0][1]) < 4e-3


@pytest.mark.parametrize("n,cin,h,cout,k,s", [(2, 256, 34, 256, 3, 1), (3, 64, 21, 512, 3, 1), (5, 128, 18, 256, 4, 2), (1, 256, 18, 256, 3, 1),
                                              (5, 256, 13, 768, 3, 1)])
def test_conv_pair_kernel_matches_torch_and_single_cta(n, cin, h, cout, k, s):
    """CTA-pair (tcgen05 cta_group::2, M = 256) variant of the N = 256 convolution tile: forward with the fused epilogue and dgrad
    against torch and against the single-CTA kernel; odd tile counts (the second CTA of the last pair has no pixels) included."""
    from de_i2i_gan_b200 import ops
    g = torch.Generator().manual_seed(7 * n + cin)
    x = q(torch.randn(n, cin, h, h, generator=g))
    wt = q(torch.randn(cout, cin, k, k, generator=g) / (cin * k * k) ** 0.5)
    b = torch.randn(cout, generator=g)
    spec = ops.ConvSpec(cin, cout, k, s, 0, "zeros")
    w_fwd, w_dg = ops.pack_weight(wt.cuda(), spec)
    xr = x.clone().requires_grad_(True)
    yr = F.leaky_relu(F.conv2d(xr, wt, b, stride=s), 0.2)
    pre = F.conv2d(x, wt, None, stride=s)
    dy = q(torch.randn(pre.shape, generator=g))
    (dx_ref,) = torch.autograd.grad(F.conv2d(xr, wt, None, stride=s), xr, dy)
    outs = {}
    for pair in (1, 0):
        ops.set_tuning("tma_pair", pair)
        try:
            y = ops.conv2d_fwd(to_dev(x), w_fwd, spec, bias=b.cuda(), act="leaky_relu")
            k_f = ops.last_kernel()
            dx = ops.conv2d_dgrad(to_dev(dy), w_dg, spec, (h, h)) if cin % 256 == 0 and s == 1 else None
            k_d = ops.last_kernel()
        finally:
            ops.set_tuning("tma_pair", 3)
        torch.cuda.synchronize()
        assert ("pair" in k_f) == bool(pair), k_f
        outs[pair] = (from_dev(y, cout), from_dev(dx, cin) if dx is not None else None)
        if dx is not None:
            assert ("pair" in k_d) == bool(pair), k_d
    e_f = rel_err(outs[1][0], yr.detach())
    assert e_f < TOL_BF16, e_f
    assert torch.equal(outs[1][0], outs[0][0])
    if outs[1][1] is not None:
        assert rel_err(outs[1][1], dx_ref) < TOL_BF16
        assert torch.equal(outs[1][1], outs[0][1])
