"""torch.autograd.Function wrappers that put the CUDA kernels behind PyTorch's autograd graph.

PyTorch only provides the graph, device memory and streams; every forward and backward body calls libdei2i.so.
Activations travel between blocks as ``Nhwc`` carriers (BF16 [N,H,W,C_ld] + real channel count); NCHW FP32 exists
only at the model edge (images, logits, losses), exactly where the reference's L2 code touches tensors.
"""
from contextlib import contextmanager
from typing import Optional

import os

import torch
from torch.autograd import Function

from . import ops
from .ops import ConvSpec


class Nhwc:
    """BF16 NHWC activation carrier. ``t``: [N,H,W,C_ld] (C_ld % 8 == 0), ``c``: real channels."""
    __slots__ = ("t", "c", "cls5", "pad", "key", "stats")

    def __init__(self, t: torch.Tensor, c: int):
        self.t, self.c, self.cls5 = t, c, None
        self.stats = None  # ('batch' | 'sample', fp32 [groups,2,C_ld]) (mean, var) of ``t`` emitted by the producing convolution's epilogue
        self.key = None  # identity of the tensor a label map was made from (SPADE caches gamma|beta per label tensor)
        self.pad = 0  # pixels of reflect padding already present on every side of ``t`` (written by the producer for its consumer)

    @property
    def shape(self):  # logical NCHW shape, like the reference's tensors
        n, h, w, _ = self.t.shape
        return (n, self.c, h - 2 * self.pad, w - 2 * self.pad)


# ------------------------------------------------------------------------------------------------ weight preparation
_cache = None  # dict while inside weight_cache(), else None


@contextmanager
def weight_cache():
    """Inside this scope packed BF16 weights are built once per parameter tensor (weights do not change within one
    loss computation: 4 G + 2 D passes of a G-step share them).  Spectral-norm state is never cached."""
    global _cache
    outer = _cache
    if outer is None:
        _cache = {}
    try:
        yield
    finally:
        if outer is None:
            _cache = None


BATCH_LABEL_SETS = os.environ.get("DEI2I_BATCH_LABEL_SETS", "1") != "0"  # A/B switch of announce_label_sets


def label_key(labels: torch.Tensor):
    """Identity of a label tensor inside one weight_cache() scope (the generator stamps it on the carrier it hands to SPADE)."""
    return (labels.data_ptr(), labels._version, tuple(labels.shape))


def announce_label_sets(*label_tensors):
    """The loss graphs of a training step condition the generator on TWO label tensors (defect labels, normal labels:
    defectgan_model.py:185-205).  Announced at the top of a weight_cache() scope, every SPADE module evaluates its class-map
    convolutions for both in ONE batched pass (models/networks/normalization.py) - half the launches of kernels that are
    launch-latency bound (32 x 25 pixels each).  No effect outside a scope or when fewer than two tensors are given."""
    if _cache is not None and len(label_tensors) > 1 and BATCH_LABEL_SETS:
        _cache["label_sets"] = tuple(label_tensors)


class Weights:
    __slots__ = ("w_fwd", "w_dgrad", "scale", "sn")

    def __init__(self, w_fwd, w_dgrad, scale, sn):
        self.w_fwd, self.w_dgrad, self.scale, self.sn = w_fwd, w_dgrad, scale, sn


def _pack(w: torch.Tensor, spec: ConvSpec):
    key = (w.data_ptr(), tuple(w.shape), spec)
    _record_pack(w, None, spec)
    if _cache is not None and key in _cache:
        return _cache[key]
    packed = ops.pack_weight(w.detach().contiguous(), spec, None, True)
    if _cache is not None:
        _cache[key] = packed
    return packed


# Packing every weight of a network in one launch (ops.pack_weight_multi) at the top of its forward instead of two ~8 us launches
# per convolution (168 per D-step + G-step).  Which (parameter, spec) pairs a forward packs is recorded during the network's first
# forward (which takes the per-weight path); entries hold the Parameter OBJECTS, so re-flattened / moved storage is followed.
PACK_MULTI = os.environ.get("DEI2I_PACK_MULTI", "1") != "0"
_pack_recording = {}  # id(net) -> {(id(w), id(w2), spec): (w, w2, spec)}


def _record_pack(w, w2, spec):
    if _pack_recording and isinstance(w, torch.nn.Parameter) and (w2 is None or isinstance(w2, torch.nn.Parameter)):
        for rec in _pack_recording.values():
            rec.setdefault((id(w), id(w2), spec), (w, w2, spec))


def _pack_key(w, w2, spec):
    return (w.data_ptr(), tuple(w.shape), spec) if w2 is None else (w.data_ptr(), w2.data_ptr(), spec)


def prefetch_packs(net: torch.nn.Module):
    """See PACK_MULTI.  Only inside a weight_cache() scope (outside it nothing outlives the call that packs) and in BF16 mode."""
    if not PACK_MULTI or _cache is None or ops.ACT_DTYPE != torch.bfloat16:
        return
    plan = getattr(net, "_pack_plan", None)
    if plan is None:
        rec = _pack_recording.get(id(net))
        if rec is None:  # first forward inside a scope: record only
            _pack_recording[id(net)] = {}
            return
        plan = net._pack_plan = list(rec.values())
        del _pack_recording[id(net)]
    todo = [(w, w2, spec) for w, w2, spec in plan if _pack_key(w, w2, spec) not in _cache]
    if not todo:
        return
    with torch.no_grad():
        packed = ops.pack_weight_multi([(w.detach(), w2.detach() if w2 is not None else None, spec) for w, w2, spec in todo])
    for (w, w2, spec), pk in zip(todo, packed):
        _cache[_pack_key(w, w2, spec)] = pk


SN_MULTI = os.environ.get("DEI2I_SN_MULTI", "1") != "0"
_sn_ready = {}  # id(conv) -> (inv_sigma, u_snapshot, v_snapshot, training) produced by prefetch_spectral_norm, consumed once


_sn_recording = {}  # id(net) -> set of id(conv) seen by prepare_conv / rowpack_conv since that net's first forward


def prefetch_spectral_norm(net: torch.nn.Module, training: bool):
    """Power iteration of the spectral-normalised convolutions of `net` in four launches (ops.spectral_norm_multi) at the top
    of its forward, instead of four launches + two copies per layer.  Same per-forward semantics as torch.nn.utils.spectral_norm
    (architecture.py:68-71): exactly the wrapped convolutions that a forward call runs are iterated - which ones those are
    (the learned shortcuts of some SPADE blocks exist but never run) is recorded during the net's first forward, which still
    takes the per-layer path."""
    if not SN_MULTI:
        return
    used = getattr(net, "_sn_used", None)
    if used is None:
        rec = _sn_recording.get(id(net))
        if rec is None:  # first forward: record only
            _sn_recording[id(net)] = set()
            return
        used = [m for m in net.modules() if hasattr(m, "weight_orig") and hasattr(m, "weight_u") and id(m) in rec]
        net._sn_used = used
        del _sn_recording[id(net)]
    if not used:
        return
    with torch.no_grad():
        res = ops.spectral_norm_multi([(m.weight_orig.detach(), m.weight_u, m.weight_v) for m in used], training)
    for m, (inv, us, vs) in zip(used, res):
        _sn_ready[id(m)] = (inv, us, vs, training)


def _sn_lookup(conv, training):
    for rec in _sn_recording.values():
        rec.add(id(conv))
    ready = _sn_ready.pop(id(conv), None)
    return ready if ready is not None and ready[3] == training else None


def prepare_conv(conv: torch.nn.Module, spec: ConvSpec, training: bool) -> Weights:
    """Spectral norm (one power iteration when training, torch.nn.utils.spectral_norm semantics: u, v updated in place
    per forward call, sigma from the updated vectors) + packed weights of W_orig; 1/sigma is applied in the conv epilogue."""
    if hasattr(conv, "weight_orig"):
        w = conv.weight_orig
        ready = _sn_lookup(conv, training)
        if ready is not None:
            inv, sn = ready[0], (w.detach(), ready[1], ready[2], ready[0])
        else:
            with torch.no_grad():
                inv = ops.spectral_norm(w.detach(), conv.weight_u, conv.weight_v, training)
                sn = (w.detach(), conv.weight_u.clone(), conv.weight_v.clone(), inv)
        w_fwd, w_dg = _pack(w, spec)
        return Weights(w_fwd, w_dg, inv, sn)
    w_fwd, w_dg = _pack(conv.weight, spec)
    return Weights(w_fwd, w_dg, None, None)


def conv_param(conv: torch.nn.Module) -> torch.Tensor:
    return conv.weight_orig if hasattr(conv, "weight_orig") else conv.weight


# ------------------------------------------------------------------------------------------------ Functions
class _ToNhwc(Function):
    @staticmethod
    def forward(ctx, x):
        ctx.c = x.shape[1]
        return ops.nchw_to_nhwc(x.contiguous().float())

    @staticmethod
    def backward(ctx, g):
        return _FromNhwc.apply(g.contiguous(), ctx.c)  # a Function, so that R1 (grad of a grad w.r.t. the image) can pass


class _FromNhwc(Function):
    @staticmethod
    def forward(ctx, t, c):
        ctx.c_ld = t.shape[3]
        return ops.nhwc_to_nchw(t, c)

    @staticmethod
    def backward(ctx, g):
        return ops.nchw_to_nhwc(g.contiguous(), ctx.c_ld), None


def to_nhwc(x) -> Nhwc:
    if isinstance(x, Nhwc):
        return x
    return Nhwc(_ToNhwc.apply(x), x.shape[1])


def from_nhwc(a: Nhwc) -> torch.Tensor:
    return _FromNhwc.apply(a.t, a.c)


# Statistics out of the convolution epilogue (dei2i_conv_epilogue.stats_partial) for outputs of up to this many channels.  Measured on
# the 256 x 256 / batch 32 step: 0 -> 81.73 ms, 128 -> 81.78 ms, 512 -> 82.31 ms per step - the epilogue work costs what the separate
# statistics pass over the (mostly L2-resident) tensor saves, so the consumer computes them by default (dei2i_channel_stats).
# Statistics of a convolution's output for its BatchNorm / InstanceNorm consumer out of the convolution epilogue (per-tile (mean, M2)
# pairs, merged in a fixed order) instead of a separate pass over the stored tensor.  Worth it where the main loop leaves the
# epilogue warps slack - 128 and 256 output channels: 77.2 -> 76.1 ms per step - and not for the 64-channel layers at 256^2, whose
# epilogue is on the critical path (+1.0 ms) - measured on one box, tools/gpu_ab.sh.
EPI_STATS_MAX_COUT = int(os.environ.get("DEI2I_EPI_STATS_MAX_COUT", "256"))
EPI_STATS_MIN_COUT = int(os.environ.get("DEI2I_EPI_STATS_MIN_COUT", "128"))


_grad_sink = False


@contextmanager
def accumulate_into_grad(enabled=True):
    """Around a trainer's ``loss.backward()``: weight gradients of the convolutions are accumulated by the wgrad kernel itself
    (beta = 1) straight into the parameter's existing ``.grad`` buffer - FlatAdam's flat gradient vector, zeroed by zero_grad - instead
    of being returned to autograd, whose AccumulateGrad then launches one `grad += dw` kernel per arrival (318 per D-step + G-step).
    Same FP32 additions in the same order (the kernel fuses a spectral-normed weight's final `* 1/sigma` with the accumulate into one
    FMA: last-bit differences in a few elements).  Not for ``torch.autograd.grad`` and not while
    post-accumulate-grad hooks must fire (the bucketed all-reduce overlap): those leave it off and get the ordinary path."""
    global _grad_sink
    outer = _grad_sink
    _grad_sink = bool(enabled)
    try:
        yield
    finally:
        _grad_sink = outer


class _Conv(Function):
    """y = act(scale*conv(x) + bias + noise_w*noise) + residual   and its dgrad / wgrad / bias / noise-weight grads.
    want_stats: the epilogue also emits the (mean, var) statistics its BatchNorm / InstanceNorm consumer needs (second, non-
    differentiable output; None where the kernel cannot).  out_pad: y carries the reflect padding of the convolution that consumes
    it (written by the same call); the gradient then arrives for the padded tensor and is folded in the backward - fused into
    the activation backward where there is one."""

    @staticmethod
    def forward(ctx, x, w_param, bias, noise_w, residual, wts: Weights, spec: ConvSpec, act, noise, out_fp32, logical_hw,
                want_stats=None, out_pad=0):
        assert not (act and residual is not None), "activation mask cannot be recovered once a residual is added"
        assert not (out_pad and residual is not None)
        # nn.Conv2d(padding_mode='reflect') = F.pad(reflect) + valid convolution (torch's own _conv_forward): the mirrored
        # border is materialised once (it is also what wgrad re-reads), the convolution itself then runs on the TMA path
        ctx.rpad = spec.pad if (spec.pad_mode == "reflect" and spec.pad > 0 and logical_hw is None) else 0
        if ctx.rpad:
            x = ops.reflect_pad(x, ctx.rpad)
            spec = spec.unpadded()
        res = ops.conv2d_fwd(x, wts.w_fwd, spec, bias=bias, noise=noise, noise_weight=noise_w, residual=residual, act=act,
                             out_fp32=out_fp32, logical_hw=logical_hw, scale=wts.scale, want_stats=want_stats, out_pad=out_pad)
        y, stats = res if want_stats is not None else (res, None)
        ctx.wts, ctx.spec, ctx.act, ctx.noise, ctx.logical_hw, ctx.out_pad = wts, spec, act, noise, logical_hw, out_pad
        ctx.has = (bias is not None, noise_w is not None, residual is not None)
        ctx.w_param = w_param
        ctx.save_for_backward(x, y if act else None)
        if want_stats is None:
            return y
        if stats is None:
            stats = y.new_zeros(0, dtype=torch.float32)  # "not available": the consumer computes them itself
        ctx.mark_non_differentiable(stats)
        return y, stats

    @staticmethod
    def backward(ctx, dy, _dstats=None):
        x, y = ctx.saved_tensors
        wts, spec = ctx.wts, ctx.spec
        has_bias, has_noise, has_res = ctx.has
        if dy.dtype != ops.ACT_DTYPE:
            dy = dy.to(ops.ACT_DTYPE)
        dy = dy.contiguous()
        d_bias = None
        need_bias = has_bias and ctx.needs_input_grad[2]
        pad = ctx.out_pad
        if ctx.act or need_bias:
            yb = y if (y is None or y.dtype == ops.ACT_DTYPE) else y.to(ops.ACT_DTYPE)
            dz, d_bias = ops.act_bwd(yb, dy, ctx.act, want_dx=bool(ctx.act) or pad > 0, want_bias_grad=need_bias, pad=pad)
            if ctx.act or pad:
                dy = dz
            if d_bias is not None:
                d_bias = d_bias[: spec.cout]
        elif pad:
            dy = ops.reflect_fold(dy, pad)
        d_res = dy if (has_res and ctx.needs_input_grad[4]) else None
        d_nw = None
        if has_noise and ctx.needs_input_grad[3]:
            d_nw = ops.noise_wgrad(dy, ctx.noise, torch.empty(1, device=dy.device), spec.cout).reshape(1, 1, 1, 1)
        dx = None
        if ctx.needs_input_grad[0]:
            assert ctx.logical_hw is None, "no input gradient through a resized (label map) input"
            dx = ops.conv2d_dgrad(dy, wts.w_dgrad, spec, (x.shape[1], x.shape[2]), scale=wts.scale)
            if ctx.rpad:
                dx = ops.reflect_fold(dx, ctx.rpad)
        dw = None
        if ctx.needs_input_grad[1]:
            w = ctx.w_param
            sink = w.grad if (_grad_sink and w.is_leaf) else None
            if sink is not None and sink.dtype == torch.float32 and sink.is_contiguous() \
                    and tuple(sink.shape) == (spec.cout, spec.cin, spec.kh, spec.kw_):
                ops.conv2d_wgrad(x, dy, sink, spec, sn=wts.sn, logical_hw=ctx.logical_hw, accumulate=True)  # accumulate_into_grad
            else:
                dw = torch.empty(spec.cout, spec.cin, spec.kh, spec.kw_, dtype=torch.float32, device=dy.device)
                ops.conv2d_wgrad(x, dy, dw, spec, sn=wts.sn, logical_hw=ctx.logical_hw)
        return dx, dw, d_bias, d_nw, d_res, None, None, None, None, None, None, None, None


def conv(x: Nhwc, conv_mod, spec: ConvSpec, training: bool, bias=None, act=None, noise=None, noise_w=None,
         residual: Optional[Nhwc] = None, out_fp32=False, logical_hw=None, wts: Weights = None, want_stats=None, out_pad=0) -> Nhwc:
    """want_stats ('batch' | 'sample'): ask the epilogue for the statistics of the output (carrier ``.stats``); out_pad: ask it to
    write the consumer's reflect padding (carrier ``.pad``) - granted only where the TMA kernel runs the geometry."""
    wts = wts or prepare_conv(conv_mod, spec, training)
    if want_stats is not None and not (EPI_STATS_MIN_COUT <= spec.cout_ld <= EPI_STATS_MAX_COUT):
        want_stats = None
    if out_pad:
        n, h, w = x.t.shape[0], x.t.shape[1], x.t.shape[2]
        if spec.pad_mode == "reflect" and spec.pad > 0:  # _Conv pads the input first and runs the 'valid' convolution
            geo, h, w = spec.unpadded(), h + 2 * spec.pad, w + 2 * spec.pad
        else:
            geo = spec
        if out_fp32 or logical_hw is not None or residual is not None or not ops.conv_fwd_is_tma(geo, n, h, w):
            out_pad = 0
    res = _Conv.apply(x.t, conv_param(conv_mod), bias, noise_w, residual.t if residual is not None else None, wts, spec, act,
                      noise, out_fp32, logical_hw, want_stats, out_pad)
    if want_stats is None:
        out = Nhwc(res, spec.cout)
    else:
        out = Nhwc(res[0], spec.cout)
        if res[1].numel():
            out.stats = (want_stats, res[1])
    out.pad = out_pad
    return out


def rowpack_applicable(spec: ConvSpec) -> bool:
    """Few-channel wide-filter stride-1 'same' reflect convolutions (the generator stem 3->64 7x7); BF16 path only."""
    return (ops.ACT_DTYPE == torch.bfloat16 and spec.cin <= 8 and 1 < spec.kw_ <= 8 and spec.stride == 1 and spec.pad_mode == "reflect"
            and 2 * spec.pad == spec.kh - 1 and spec.kh == spec.kw_ and spec.cout_ld >= 64)


def _rowpack_weight(w: torch.Tensor, spec: ConvSpec, spec2: ConvSpec):
    """[Cout,C,kh,kw] -> [Cout,64,kh,1] with input channel s*8+c = (filter column s, image channel c); packed GEMM operands."""
    key = ("rowpack", w.data_ptr(), tuple(w.shape))
    if _cache is not None and key in _cache:
        return _cache[key]
    w2 = torch.zeros(spec.cout, 8, 8, spec.kh, dtype=torch.float32, device=w.device)
    w2[:, : spec.kw_, : spec.cin] = w.detach().permute(0, 3, 1, 2)
    packed = ops.pack_weight(w2.reshape(spec.cout, 64, spec.kh, 1), spec2, None, True)
    if _cache is not None:
        _cache[key] = packed
    return packed


class _RowpackConv(Function):
    """scale * conv(x, W) for a fp32 NCHW image x through the row-packed layout (dei2i_rowpack_fwd): the kh x kw convolution
    over C <= 8 channels runs as a kh x 1 convolution over 64 channels on the TMA / tcgen05 kernels."""

    @staticmethod
    def forward(ctx, x, w_param, wts_sn, scale, spec: ConvSpec, want_stats=None):
        spec2 = ConvSpec(64, spec.cout, spec.kh, 1, 0, "zeros", 1)
        w_fwd, w_dg = _rowpack_weight(w_param, spec, spec2)
        xp = ops.rowpack_fwd(x.contiguous().float(), spec.kw_, spec.pad)
        ctx.flops = spec.flops(x.shape[0], x.shape[2], x.shape[3])  # algorithmic FLOPs of the ORIGINAL kh x kw x C convolution
        res = ops.conv2d_fwd(xp, w_fwd, spec2, scale=scale, alg_flops=ctx.flops, want_stats=want_stats)
        ctx.spec, ctx.spec2, ctx.w_dg, ctx.scale, ctx.sn = spec, spec2, w_dg, scale, wts_sn
        ctx.save_for_backward(xp)
        if want_stats is None:
            return res
        y, stats = res
        if stats is None:
            stats = y.new_zeros(0, dtype=torch.float32)
        ctx.mark_non_differentiable(stats)
        return y, stats

    @staticmethod
    def backward(ctx, dy, _dstats=None):
        (xp,) = ctx.saved_tensors
        spec, spec2 = ctx.spec, ctx.spec2
        dy = dy.contiguous()
        dx = dw = None
        if ctx.needs_input_grad[0]:
            dxp = ops.conv2d_dgrad(dy, ctx.w_dg, spec2, (xp.shape[1], xp.shape[2]), scale=ctx.scale, alg_flops=ctx.flops)
            dx = ops.rowpack_bwd(dxp, spec.cin, spec.kw_, spec.pad)
        if ctx.needs_input_grad[1]:
            dw2 = torch.empty(spec.cout, 64, spec.kh, 1, dtype=torch.float32, device=dy.device)
            ops.conv2d_wgrad(xp, dy, dw2, spec2, alg_flops=ctx.flops)
            g = dw2.view(spec.cout, 8, 8, spec.kh)[:, : spec.kw_, : spec.cin].permute(0, 2, 3, 1)  # [Cout,C,kh,kw]: d/d(W/sigma)
            if ctx.sn is not None:
                # spectral_norm backward (u, v constants): dW_orig = (G - <G, W/sigma> u v^T) / sigma
                w0, u, v, inv = ctx.sn
                dot = (g * w0).sum() * inv[0]
                dw = (g - dot * torch.outer(u, v).view_as(w0)) * inv[0]
            else:
                dw = g.contiguous()
        return dx, dw, None, None, None, None


def rowpack_conv(x: torch.Tensor, conv_mod, spec: ConvSpec, training: bool, want_stats=None) -> Nhwc:
    """Stem convolution on a plain fp32 NCHW image (bias-free).  Spectral norm as in prepare_conv."""
    sn = scale = None
    if hasattr(conv_mod, "weight_orig"):
        w = conv_mod.weight_orig
        ready = _sn_lookup(conv_mod, training)
        if ready is not None:
            inv, sn = ready[0], (w.detach(), ready[1], ready[2], ready[0])
        else:
            with torch.no_grad():
                inv = ops.spectral_norm(w.detach(), conv_mod.weight_u, conv_mod.weight_v, training)
                sn = (w.detach(), conv_mod.weight_u.clone(), conv_mod.weight_v.clone(), inv)
        scale = inv
    else:
        w = conv_mod.weight
    if want_stats is not None and not (EPI_STATS_MIN_COUT <= spec.cout_ld <= EPI_STATS_MAX_COUT):
        want_stats = None
    res = _RowpackConv.apply(x, w, sn, scale, spec, want_stats)
    if want_stats is None:
        return Nhwc(res, spec.cout)
    out = Nhwc(res[0], spec.cout)
    if res[1].numel():
        out.stats = (want_stats, res[1])
    return out


def linear(x: torch.Tensor, lin: torch.nn.Linear, act=None) -> torch.Tensor:
    """nn.Linear on fp32 [N, Cin] as a 1x1 convolution over a [N,1,1,Cin] map (same GEMM kernels, FP32 logits out)."""
    cout, cin = lin.weight.shape
    spec = ConvSpec(cin, cout, 1)
    w4 = lin.weight.view(cout, cin, 1, 1)
    wts = Weights(*_pack(w4, spec), None, None)
    xb = x.to(ops.ACT_DTYPE)
    if spec.cin_ld != cin:
        xb = torch.nn.functional.pad(xb, (0, spec.cin_ld - cin))
    y = _Conv.apply(xb.reshape(x.shape[0], 1, 1, spec.cin_ld).contiguous(), w4, lin.bias, None, None, wts, spec, act, None, True, None)
    return y[:, 0, 0, :cout]


class _ConvCat(Function):
    """Two convolutions sharing their input fused into one GEMM with N = Cout_a + Cout_b (SPADE gamma|beta,
    normalization.py:31-32; generator heads, generator.py:268-269).  Parameters stay separate tensors."""

    @staticmethod
    def forward(ctx, x, wa, wb, ba, bb, spec: ConvSpec, out_fp32):
        w_fwd, w_dg = _pack_cat(wa, wb, spec)
        bias = torch.cat([ba.detach(), bb.detach()]) if ba is not None else None
        ctx.rpad = spec.pad if (spec.pad_mode == "reflect" and spec.pad > 0) else 0
        if ctx.rpad:
            x = ops.reflect_pad(x, ctx.rpad)
            spec = spec.unpadded()
        y = ops.conv2d_fwd(x, w_fwd, spec, bias=bias, out_fp32=out_fp32)
        ctx.spec, ctx.w_dg, ctx.ca, ctx.has_bias = spec, w_dg, wa.shape[0], ba is not None
        ctx.save_for_backward(x)
        return y

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        spec, ca = ctx.spec, ctx.ca
        if dy.dtype != ops.ACT_DTYPE:
            dy = dy.to(ops.ACT_DTYPE)
        dy = dy.contiguous()
        dba = dbb = None
        if ctx.has_bias:
            _, db = ops.act_bwd(None, dy, None, want_dx=False, want_bias_grad=True)
            dba, dbb = db[:ca], db[ca: spec.cout]
        dx = ops.conv2d_dgrad(dy, ctx.w_dg, spec, (x.shape[1], x.shape[2])) if ctx.needs_input_grad[0] else None
        if dx is not None and ctx.rpad:
            dx = ops.reflect_fold(dx, ctx.rpad)
        dw = torch.empty(spec.cout, spec.cin, spec.kh, spec.kw_, dtype=torch.float32, device=dy.device)
        ops.conv2d_wgrad(x, dy, dw, spec)
        return dx, dw[:ca], dw[ca:], dba, dbb, None, None


def _pack_cat(wa, wb, spec):
    key = (wa.data_ptr(), wb.data_ptr(), spec)
    _record_pack(wa, wb, spec)
    if _cache is not None and key in _cache:
        return _cache[key]
    packed = ops.pack_weight(torch.cat([wa.detach(), wb.detach()], 0).contiguous(), spec, None, True)
    if _cache is not None:
        _cache[key] = packed
    return packed


def conv_cat(x: Nhwc, conv_a, conv_b, spec: ConvSpec, out_fp32=False) -> Nhwc:
    y = _ConvCat.apply(x.t, conv_a.weight, conv_b.weight, conv_a.bias, conv_b.bias, spec, out_fp32)
    return Nhwc(y, spec.cout)


# How the BatchNorm backward takes the gradient of a PADDED output (the consumer's reflect padding was written by the forward):
#   0: dei2i_reflect_fold into a dense tensor first (a full read + write pass), 1: folded while read (mirror reads in both passes of the
#   backward; measured slower), 2: the border is folded IN PLACE into the interior (3 % of the pixels) and the backward reads the
#   interior through the padded pitch.
BN_FUSED_FOLD = int(os.environ.get("DEI2I_BN_FUSED_FOLD", "2"))  # same box: 75.3 -> 74.85 ms per step


class _BatchNorm(Function):
    @staticmethod
    def forward(ctx, x, gamma, beta, residual, bn, training, act, out_pad, stats=None):
        y, save = ops.bn_fwd(x, gamma.detach(), beta.detach(), bn.running_mean, bn.running_var, training, act=act,
                             residual=residual, momentum=bn.momentum if bn.momentum is not None else 0.1, eps=bn.eps,
                             out_pad=out_pad, stats=stats)
        if training and bn.num_batches_tracked is not None:
            bn.num_batches_tracked += 1
        ctx.training, ctx.act, ctx.has_res, ctx.out_pad = training, act, residual is not None, out_pad
        ctx.save_for_backward(x, gamma, beta, save)
        return y

    @staticmethod
    def backward(ctx, dy):
        x, gamma, beta, save = ctx.saved_tensors
        dy = dy.contiguous()
        pad = ctx.out_pad
        # see BN_FUSED_FOLD: folding while reading measured slower than fold + dense backward (592 vs 373 us on the 256 x 256 x 64 layer);
        # the in-place border fold is the default.  The residual branch needs the folded dense tensor anyway.
        if pad and (ctx.has_res or not BN_FUSED_FOLD):
            dy, pad = ops.reflect_fold(dy, pad), 0
        # the consumer's gradient covers the padded frame: bn_bwd folds the mirrored border back while reading it
        dx, dg, db = ops.bn_bwd(x, dy, gamma.detach(), beta.detach(), save, ctx.training, act=ctx.act, dy_pad=pad,
                                fold_in_place=BN_FUSED_FOLD == 2)
        return dx, dg, db, (dy if ctx.has_res else None), None, None, None, None, None


def batch_norm(x: Nhwc, bn: torch.nn.BatchNorm2d, act=None, residual: Optional[Nhwc] = None, out_pad: int = 0) -> Nhwc:
    """out_pad > 0: the result already carries the reflect padding of the convolution that consumes it (Nhwc.pad)."""
    training = bn.training or not bn.track_running_stats
    stats = x.stats[1] if (training and x.stats is not None and x.stats[0] == "batch") else None
    y = _BatchNorm.apply(x.t, bn.weight, bn.bias, residual.t if residual is not None else None, bn, training, act, out_pad, stats)
    out = Nhwc(y, x.c)
    out.pad = out_pad
    return out


class _Spade(Function):
    @staticmethod
    def forward(ctx, x, gb, up, relu, eps, stats=None):
        out, mr = ops.spade_fwd(x, gb, up, relu, eps, stats=stats)
        ctx.up, ctx.relu = up, relu
        ctx.save_for_backward(x, gb, out, mr)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, gb, out, mr = ctx.saved_tensors
        dgb, dx = ops.spade_bwd(x, gb, out, dout.contiguous(), mr, ctx.up, ctx.relu)
        return dx, dgb, None, None, None, None


def spade_modulate(x: Nhwc, gb: Nhwc, up=1, relu=True, eps=1e-5) -> Nhwc:
    return Nhwc(_Spade.apply(x.t, gb.t, up, relu, eps, _sample_stats(x)), x.c)


def _sample_stats(x: Nhwc):
    """Per-image statistics the producing convolution already emitted for this carrier, else None (computed by the consumer)."""
    return x.stats[1] if (x.stats is not None and x.stats[0] == "sample") else None


class _SpadeCls(Function):
    """SPADE modulation with gamma|beta on the [N,5,5,2C] border-class map of a constant label map."""

    @staticmethod
    def forward(ctx, x, gb5, up, relu, eps, out_pad, stats=None):
        out, mr = ops.spade_cls_fwd(x, gb5, up, relu, eps, out_pad, stats=stats)
        ctx.up, ctx.relu, ctx.out_pad = up, relu, out_pad
        ctx.save_for_backward(x, gb5, out, mr)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, gb5, out, mr = ctx.saved_tensors
        dgb5, dx = ops.spade_cls_bwd(x, gb5, out, dout.contiguous(), mr, ctx.up, ctx.relu, ctx.out_pad)
        return dx, dgb5.to(ops.ACT_DTYPE), None, None, None, None, None


def spade_modulate_cls(x: Nhwc, gb5: Nhwc, up=1, relu=True, eps=1e-5, out_pad=0) -> Nhwc:
    return Nhwc(_SpadeCls.apply(x.t, gb5.t, up, relu, eps, out_pad, _sample_stats(x)), x.c)


class _UpsamplePad(Function):
    @staticmethod
    def forward(ctx, x, up, pad):
        ctx.up, ctx.pad, ctx.hw = up, pad, (x.shape[1], x.shape[2])
        return ops.upsample_pad(x, up, pad)

    @staticmethod
    def backward(ctx, dout):
        return ops.upsample_pad_bwd(dout.contiguous(), ctx.hw, ctx.up, ctx.pad), None, None


def upsample_pad(x: Nhwc, up=1, pad=(0, 0, 0, 0)) -> Nhwc:
    """nn.Upsample(scale_factor=up) followed by a zero frame (top, left, bottom, right)."""
    return Nhwc(_UpsamplePad.apply(x.t, up, tuple(pad)), x.c)


class _MaxPool(Function):
    @staticmethod
    def forward(ctx, x, k, stride, pad):
        out, idx = ops.maxpool_fwd(x, k, stride, pad)
        ctx.cfg, ctx.hw = (k, stride, pad), (x.shape[1], x.shape[2])
        ctx.save_for_backward(idx)
        return out

    @staticmethod
    def backward(ctx, dout):
        (idx,) = ctx.saved_tensors
        return ops.maxpool_bwd(dout.contiguous(), idx, ctx.hw, *ctx.cfg), None, None, None


def max_pool(x: Nhwc, k, stride, pad) -> Nhwc:
    return Nhwc(_MaxPool.apply(x.t, k, stride, pad), x.c)


class _AvgPool(Function):
    @staticmethod
    def forward(ctx, x, k, scale):
        ctx.k, ctx.scale = k, scale
        return ops.avgpool_fwd(x, k, scale)

    @staticmethod
    def backward(ctx, dout):
        return _AvgPoolBwd.apply(dout, ctx.k, ctx.scale), None, None


class _AvgPoolBwd(Function):
    @staticmethod
    def forward(ctx, dout, k, scale):
        ctx.k, ctx.scale = k, scale
        return ops.avgpool_bwd(dout.contiguous(), k, scale)

    @staticmethod
    def backward(ctx, ddx):
        return _AvgPool.apply(ddx.contiguous(), ctx.k, ctx.scale), None, None


def avg_pool(x: Nhwc, k, scale=1.0) -> Nhwc:
    return Nhwc(_AvgPool.apply(x.t, k, scale), x.c)


class _Gap(Function):
    @staticmethod
    def forward(ctx, x, c):
        ctx.hw, ctx.c_ld = (x.shape[1], x.shape[2]), x.shape[3]
        return ops.gap_fwd(x)[:, :c]

    @staticmethod
    def backward(ctx, dout):
        g = dout.contiguous().float()
        if g.shape[1] != ctx.c_ld:
            g = torch.nn.functional.pad(g, (0, ctx.c_ld - g.shape[1]))
        return ops.gap_bwd(g, ctx.hw), None


def global_avg_pool(x: Nhwc) -> torch.Tensor:
    """nn.AdaptiveAvgPool2d(1) + flatten: fp32 [N, C]."""
    return _Gap.apply(x.t, x.c)


# ---------------------------------------------------------------------------------------------- twice-differentiable set
# StarGAN-v2's R1 regulariser (stargan-v2/core/solver.py:573-583) differentiates the discriminator's input gradient again
# (create_graph=True).  The discriminator is piecewise linear (convolutions, LeakyReLU, average pooling, adds), so the backward
# of every op below is itself expressed with Functions: the engine can record it and differentiate it once more.
class _ActBwd(Function):
    """dz = dy * act'(y) for relu / leaky_relu (derivative piecewise constant in y, linear in dy: its own adjoint)."""

    @staticmethod
    def forward(ctx, dy, y, act):
        ctx.act = act
        ctx.save_for_backward(y)
        return ops.act_bwd(y, dy.contiguous(), act)[0]

    @staticmethod
    def backward(ctx, ddz):
        (y,) = ctx.saved_tensors
        return _ActBwd.apply(ddz, y, ctx.act), None, None


class _Act(Function):
    @staticmethod
    def forward(ctx, x, act):
        assert act in ("relu", "leaky_relu")
        y = ops.act_fwd(x, act)
        ctx.act = act
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        return _ActBwd.apply(dy, y, ctx.act), None


def activation(x: Nhwc, act="leaky_relu") -> Nhwc:
    return Nhwc(_Act.apply(x.t, act), x.c)


class _Axpby(Function):
    """sa * a + sb * b."""

    @staticmethod
    def forward(ctx, a, b, sa, sb):
        ctx.sa, ctx.sb, ctx.has_b = sa, sb, b is not None
        return ops.axpby(a.contiguous(), b.contiguous() if b is not None else None, sa, sb)

    @staticmethod
    def backward(ctx, g):
        ga = _Axpby.apply(g, None, ctx.sa, 0.0) if ctx.needs_input_grad[0] else None
        gb = None
        if ctx.has_b and ctx.needs_input_grad[1]:
            gb = ga if (ga is not None and ctx.sa == ctx.sb) else _Axpby.apply(g, None, ctx.sb, 0.0)
        return ga, gb, None, None


def add_scaled(a: Nhwc, b: Nhwc, s: float) -> Nhwc:
    """(a + b) * s - the StarGAN-v2 residual merge (x + r) / sqrt(2)."""
    return Nhwc(_Axpby.apply(a.t, b.t, s, s), a.c)


_input_grad_only = False


@contextmanager
def input_grad_only():
    """Inside this scope _ConvL.backward computes only the input gradient (torch.autograd.grad(..., inputs=image) of R1)."""
    global _input_grad_only
    old, _input_grad_only = _input_grad_only, True
    try:
        yield
    finally:
        _input_grad_only = old


class _DgradL(Function):
    """dx = dgrad(dy, W); differentiable in dy (-> a forward convolution of the incoming gradient) and in W (-> a wgrad)."""

    @staticmethod
    def forward(ctx, dy, w, spec, in_hw):
        ctx.spec = spec
        ctx.save_for_backward(dy, w)
        return ops.conv2d_dgrad(dy.contiguous(), _pack(w, spec)[1], spec, in_hw)

    @staticmethod
    def backward(ctx, ddx):
        dy, w = ctx.saved_tensors
        spec = ctx.spec
        ddx = ddx.contiguous()
        d_dy = ops.conv2d_fwd(ddx, _pack(w, spec)[0], spec) if ctx.needs_input_grad[0] else None
        d_w = None
        if ctx.needs_input_grad[1]:
            d_w = torch.empty(spec.cout, spec.cin, spec.kh, spec.kw_, dtype=torch.float32, device=ddx.device)
            ops.conv2d_wgrad(ddx, dy.contiguous(), d_w, spec)
        return d_dy, d_w, None, None


class _ConvL(Function):
    """y = conv(x, W) + bias ('zeros' padding, no activation): the plain nn.Conv2d of StarGAN-v2, twice differentiable along
    the input-gradient path (dx is produced by _DgradL)."""

    @staticmethod
    def forward(ctx, x, w, bias, spec, out_fp32):
        ctx.spec, ctx.has_bias = spec, bias is not None
        ctx.save_for_backward(x, w)
        return ops.conv2d_fwd(x, _pack(w, spec)[0], spec, bias=bias.detach() if bias is not None else None, out_fp32=out_fp32)

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        spec = ctx.spec
        if dy.dtype != ops.ACT_DTYPE:
            dy = dy.to(ops.ACT_DTYPE)
        dx = _DgradL.apply(dy, w, spec, (x.shape[1], x.shape[2])) if ctx.needs_input_grad[0] else None
        dw = db = None
        if _input_grad_only:
            return dx, None, None, None, None
        dyc = dy.detach().contiguous()
        if ctx.needs_input_grad[1]:
            dw = torch.empty(spec.cout, spec.cin, spec.kh, spec.kw_, dtype=torch.float32, device=dy.device)
            ops.conv2d_wgrad(x, dyc, dw, spec)
        if ctx.has_bias and ctx.needs_input_grad[2]:
            db = ops.act_bwd(None, dyc, None, want_dx=False, want_bias_grad=True)[1][: spec.cout]
        return dx, dw, db, None, None


def conv_linear(x: Nhwc, conv_mod, out_fp32=False) -> Nhwc:
    """nn.Conv2d(k, stride 1, zero padding, optional bias) on the twice-differentiable path."""
    kh, kw = conv_mod.kernel_size
    pad = conv_mod.padding[0] if isinstance(conv_mod.padding, (tuple, list)) else conv_mod.padding
    assert conv_mod.padding_mode == "zeros" and conv_mod.stride[0] == 1 and kh == kw
    spec = ConvSpec(conv_mod.in_channels, conv_mod.out_channels, kh, 1, pad, "zeros")
    return Nhwc(_ConvL.apply(x.t, conv_mod.weight, conv_mod.bias, spec, out_fp32), spec.cout)


class _AdainAct(Function):
    @staticmethod
    def forward(ctx, x, scale, shift, slope, eps):
        scale = scale.contiguous().float()
        out, mr = ops.adain_act_fwd(x, scale, shift.contiguous().float(), slope, eps)
        ctx.slope = slope
        ctx.save_for_backward(x, out, scale, mr)
        return out

    @staticmethod
    def backward(ctx, dout):
        x, out, scale, mr = ctx.saved_tensors
        dx, dscale, dshift = ops.adain_act_bwd(x, out, dout.contiguous(), scale, mr, ctx.slope)
        return dx, dscale, dshift, None, None


def adain_act(x: Nhwc, scale: torch.Tensor, shift: torch.Tensor, slope=0.2, eps=1e-5) -> Nhwc:
    """lrelu_slope(IN(x) * scale[n,c] + shift[n,c]); scale / shift fp32 [N,C] (AdaIN: 1 + gamma, beta; affine IN: weight, bias)."""
    c_ld = x.t.shape[3]
    if scale.shape[1] != c_ld:  # carrier channels are padded to a multiple of 8: the padding lanes stay zero (scale = shift = 0)
        scale = torch.nn.functional.pad(scale, (0, c_ld - scale.shape[1]))
        shift = torch.nn.functional.pad(shift, (0, c_ld - shift.shape[1]))
    return Nhwc(_AdainAct.apply(x.t, scale, shift, slope, eps), x.c)


class _Compose(Function):
    @staticmethod
    def forward(ctx, hd, x, cycle_gan):
        out, prob, _ = ops.compose_fwd(hd, x, cycle_gan, want_nhwc=False)
        ctx.cycle = cycle_gan
        ctx.save_for_backward(hd, x)
        return out, prob

    @staticmethod
    def backward(ctx, dout, dprob):
        hd, x = ctx.saved_tensors
        dhd, dx = ops.compose_bwd(hd, x, dout.contiguous() if dout is not None else None,
                                  dprob.contiguous() if dprob is not None else None, ctx.cycle, ctx.needs_input_grad[1])
        return dhd.float(), dx, None


def compose(hd: torch.Tensor, x: torch.Tensor, cycle_gan: bool):
    return _Compose.apply(hd, x.contiguous().float(), cycle_gan)


class _BceLogits(Function):
    @staticmethod
    def forward(ctx, logits, targets):
        z = logits.contiguous().float()
        if isinstance(targets, (int, float)):
            loss, grad = ops.bce_logits(z, None, float(targets))
        else:
            loss, grad = ops.bce_logits(z, targets.contiguous().float())
        ctx.save_for_backward(grad)
        return loss

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return grad * g, None


class _L1(Function):
    @staticmethod
    def forward(ctx, a, b):
        a = a.contiguous().float()
        ga = torch.empty_like(a) if ctx.needs_input_grad[0] else None
        gb = torch.empty_like(a) if (b is not None and ctx.needs_input_grad[1]) else None
        loss = ops.l1(a, b.contiguous().float() if b is not None else None, ga, gb)
        ctx.save_for_backward(ga, gb)
        return loss

    @staticmethod
    def backward(ctx, g):
        ga, gb = ctx.saved_tensors
        return (ga * g if ga is not None else None), (gb * g if gb is not None else None)


class _Mse(Function):
    @staticmethod
    def forward(ctx, a, b):
        a = a.contiguous().float()
        ga = torch.empty_like(a) if ctx.needs_input_grad[0] else None
        gb = torch.empty_like(a) if (b is not None and ctx.needs_input_grad[1]) else None
        loss = ops.mse(a, b.contiguous().float() if b is not None else None, ga, gb)
        ctx.save_for_backward(ga, gb)
        return loss

    @staticmethod
    def backward(ctx, g):
        ga, gb = ctx.saved_tensors
        return (ga * g if ga is not None else None), (gb * g if gb is not None else None)


class _CeSoft(Function):
    @staticmethod
    def forward(ctx, logits, targets):
        loss, grad = ops.ce_soft(logits.contiguous().float(), targets.contiguous().float())
        ctx.save_for_backward(grad)
        return loss

    @staticmethod
    def backward(ctx, g):
        (grad,) = ctx.saved_tensors
        return grad * g, None


def mse_loss(a, b=None):
    return _Mse.apply(a, b)


def cross_entropy_soft(logits, targets):
    """F.cross_entropy(logits [B,C], class-probability targets [B,C]) (mean over the batch)."""
    if targets.dtype in (torch.long, torch.int):
        targets = torch.nn.functional.one_hot(targets, logits.shape[1]).float()
    return _CeSoft.apply(logits, targets)


def bce_with_logits(logits, targets):
    """binary_cross_entropy_with_logits(mean). ``targets``: tensor or python scalar (all-ones / all-zeros labels)."""
    return _BceLogits.apply(logits, targets)


def l1_loss(a, b=None):
    return _L1.apply(a, b)
