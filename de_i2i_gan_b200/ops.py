"""Thin torch-tensor wrappers over the C ABI (include/dei2i.h).  PyTorch is plumbing here: device memory,
streams.  Activations are BF16 tensors of shape [N, H, W, C] (NHWC, contiguous, C a multiple of 8)."""
import ctypes as C
import os
from dataclasses import dataclass

import torch

from . import _lib
from ._lib import ConvEpilogue, ConvGeom, check

ACT = {None: 0, "none": 0, "leaky_relu": 1, "relu": 2, "tanh": 3, "sigmoid": 4}

# Activation storage type of the path.  bf16 (default): the production kernels (tcgen05 / TMA, BF16 storage, FP32 accumulation).
# fp32 (DEI2I_KERNELS=fp32 or set_precision('fp32')): the FP32-parity path of csrc/fp32_path.cu - exact FP32 arithmetic on FP32 NHWC
# activations, for comparing with the FP32 reference at the north-star's 1e-3 tolerance.  Every wrapper below dispatches on the dtype
# of the activation it is given; ACT_DTYPE decides what the model-edge converters and the weight packer produce.
#   tf32: FP32 storage as above, forward / dgrad of the TMA-able convolutions on the kind::tf32 tcgen05 kernel (10-bit mantissa
#   products, FP32 accumulation): the fast FP32-storage mode; weight gradients and everything else stay exact FP32.
_env_tuned = False  # DEI2I_TUNE applied (ensure_env_tuning)
_KERNELS = os.environ.get("DEI2I_KERNELS", "bf16").lower()
ACT_DTYPE = torch.float32 if _KERNELS in ("fp32", "f32", "tf32") else torch.bfloat16
PRECISION = "bf16"


def set_precision(name: str):
    """'bf16' | 'fp32' | 'tf32' - returns the previous name.  Switch only between loss computations (packed weights are cached per scope)."""
    global ACT_DTYPE, PRECISION
    name = {"f32": "fp32"}.get(name.lower(), name.lower())
    old = PRECISION
    ACT_DTYPE = {"bf16": torch.bfloat16, "fp32": torch.float32, "tf32": torch.float32}[name]
    PRECISION = name
    _lib.load().dei2i_set_tuning(b"f32_tf32", 1 if name == "tf32" else 0)
    return old


if _KERNELS in ("fp32", "f32", "tf32"):
    PRECISION = "tf32" if _KERNELS == "tf32" else "fp32"
    if PRECISION == "tf32":
        _lib.load().dei2i_set_tuning(b"f32_tf32", 1)


def is_f32(t) -> bool:
    return t.dtype == torch.float32
PAD = {"zeros": 0, "reflect": 1}


# Optional in-band kernel timing (bench.py's roofline): when PROFILE is a list every convolution launch is bracketed
# by CUDA events on the launching stream and appended as (kind, algorithmic_flops, start_event, end_event).
PROFILE = None
# Optional kernel-variant log (parity tests): when a list, every convolution call appends (op, geometry tag, dei2i_last_kernel()).
KERNEL_LOG = None


def _log_kernel(op, tag):
    if KERNEL_LOG is not None:
        KERNEL_LOG.append((op, tag, last_kernel()))


class _Timed:
    __slots__ = ("kind", "flops", "e0", "tag")

    def __init__(self, kind, flops, tag=""):
        self.kind, self.flops, self.tag = kind, flops, tag

    def __enter__(self):
        if PROFILE is not None:
            self.e0 = torch.cuda.Event(enable_timing=True)
            self.e0.record()

    def __exit__(self, *exc):
        if PROFILE is not None:
            e1 = torch.cuda.Event(enable_timing=True)
            e1.record()
            PROFILE.append((self.kind, self.flops, self.e0, e1, self.tag))
        return False


def pad8(c: int) -> int:
    return (c + 7) // 8 * 8


def _stream():
    if not _env_tuned:
        ensure_env_tuning()
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _p(t):
    if t is None:
        return None
    return C.c_void_p(t.data_ptr())


def _chk_act(t: torch.Tensor, name: str):
    if not (t.is_cuda and t.dtype in (torch.bfloat16, torch.float32) and t.is_contiguous() and t.dim() == 4 and t.shape[3] % 8 == 0):
        raise ValueError(f"{name}: expected a contiguous CUDA bf16 (or fp32-path) [N,H,W,C%8==0] tensor, got {t.dtype} {tuple(t.shape)}")


def _chk_f32(t: torch.Tensor, name: str):
    if not (t.is_cuda and t.dtype == torch.float32 and t.is_contiguous()):
        raise ValueError(f"{name}: expected a contiguous CUDA fp32 tensor")


@dataclass(frozen=True)
class ConvSpec:
    """Static description of one nn.Conv2d (reference: models/networks/architecture.py:95-100 etc.)."""
    cin: int        # real input channels
    cout: int       # real output channels
    k: int
    stride: int = 1
    pad: int = 0
    pad_mode: str = "zeros"
    kw: int = 0     # 0 -> square

    @property
    def kh(self):
        return self.k

    @property
    def kw_(self):
        return self.kw or self.k

    @property
    def cin_ld(self):
        return pad8(self.cin)

    @property
    def cout_ld(self):
        return pad8(self.cout)

    def out_hw(self, h, w):
        return (h + 2 * self.pad - self.kh) // self.stride + 1, (w + 2 * self.pad - self.kw_) // self.stride + 1

    def geom(self, n, h, w, hs=None, ws=None) -> ConvGeom:
        return ConvGeom(n, h, w, hs or h, ws or w, self.cin_ld, self.cout, self.cout_ld, self.kh, self.kw_, self.stride,
                        self.pad, PAD[self.pad_mode])

    def unpadded(self) -> "ConvSpec":
        """The same convolution over an input that already carries its (reflect) padding: pad = 0, 'zeros'."""
        return ConvSpec(self.cin, self.cout, self.k, self.stride, 0, "zeros", self.kw)

    def tag(self, n, h, w) -> str:
        return f"N{n} {self.cin}->{self.cout} {self.kh}x{self.kw_} s{self.stride} p{self.pad}{self.pad_mode[0]} in{h}x{w}"

    def flops(self, n, ho, wo) -> int:
        """Algorithmic FLOPs (2 * MAC over REAL channels) of one forward / dgrad / wgrad pass."""
        return 2 * n * ho * wo * self.cout * self.cin * self.kh * self.kw_

    @property
    def klass(self):
        """'dense': both GEMM dimensions >= 64 channels -> tcgen05 kernels, tensor-bound; 'skinny': HBM-bound kernels."""
        return "dense" if (self.cin_ld >= 64 and self.cout_ld >= 64) else "skinny"

    @property
    def kf(self):
        return (self.kh * self.kw_ * self.cin_ld + 63) // 64 * 64

    @property
    def kd(self):
        return (self.kh * self.kw_ * self.cout_ld + 63) // 64 * 64


def pack_weight(w: torch.Tensor, spec: ConvSpec, inv_sigma: torch.Tensor = None, need_dgrad: bool = True):
    """fp32 [Cout,Cin,kh,kw] (optionally scaled by the device scalar inv_sigma) -> packed bf16 GEMM operands."""
    _chk_f32(w, "w")
    assert tuple(w.shape) == (spec.cout, spec.cin, spec.kh, spec.kw_), (w.shape, spec)
    w_fwd = torch.empty(spec.cout, spec.kf, dtype=ACT_DTYPE, device=w.device)
    w_dg = torch.empty(spec.cin_ld, spec.kd, dtype=ACT_DTYPE, device=w.device) if need_dgrad else None
    lib = _lib.load()
    if ACT_DTYPE == torch.float32:
        assert inv_sigma is None
        check(lib.dei2i_pack_weight_f32(_p(w), spec.cout, spec.cin, spec.kh, spec.kw_, spec.cin_ld, spec.cout_ld, _p(w_fwd), _p(w_dg),
                                        _stream()), "dei2i_pack_weight_f32")
        return w_fwd, w_dg
    check(lib.dei2i_pack_weight(_p(w), _p(inv_sigma), spec.cout, spec.cin, spec.kh, spec.kw_, spec.cin_ld, spec.cout_ld,
                                _p(w_fwd), _p(w_dg), _stream()), "dei2i_pack_weight")
    return w_fwd, w_dg


def pack_weight_multi(entries):
    """entries: [(w, w2 | None, spec)] - `spec.cout` rows in total, the first `w.shape[0]` from w, the rest from w2.  All packed BF16
    matrices (forward and dgrad layout) of the list in one launch per 32 entries, carved out of ONE allocation.
    -> [(w_fwd, w_dgrad)].  BF16 mode only (the FP32-parity path packs per weight)."""
    assert ACT_DTYPE == torch.bfloat16
    if not entries:
        return []
    dev = entries[0][0].device
    al = lambda k: (k + 127) // 128 * 128  # noqa: E731  256-byte aligned matrices (TMA global addresses need 16)
    sizes = [(spec.cout * spec.kf, spec.cin_ld * spec.kd) for _, _, spec in entries]
    flat = torch.empty(sum(al(a) + al(b) for a, b in sizes), dtype=torch.bfloat16, device=dev)
    items = (_lib.PackItem * len(entries))()
    off, res = 0, []
    for i, ((w, w2, spec), (nf, nd)) in enumerate(zip(entries, sizes)):
        _chk_f32(w, "w")
        rows = w.shape[0] + (w2.shape[0] if w2 is not None else 0)
        assert rows == spec.cout and tuple(w.shape[1:]) == (spec.cin, spec.kh, spec.kw_) and w.is_contiguous(), (w.shape, spec)
        if w2 is not None:
            _chk_f32(w2, "w2")
            assert tuple(w2.shape[1:]) == tuple(w.shape[1:]) and w2.is_contiguous()
        w_fwd = flat[off:off + nf].view(spec.cout, spec.kf)
        off += al(nf)
        w_dg = flat[off:off + nd].view(spec.cin_ld, spec.kd)
        off += al(nd)
        it = items[i]
        it.w, it.w2, it.w_fwd, it.w_dgrad = _p(w), _p(w2), _p(w_fwd), _p(w_dg)
        it.rows_w, it.Cout, it.Cin_real, it.kh, it.kw, it.Cin, it.Cout_ld = w.shape[0], spec.cout, spec.cin, spec.kh, spec.kw_, spec.cin_ld, spec.cout_ld
        res.append((w_fwd, w_dg))
    check(_lib.load().dei2i_pack_weight_multi(items, len(entries), _stream()), "dei2i_pack_weight_multi")
    return res


def _kind(op, spec, flops):
    """Timing class of a convolution launch: 'dense' (tcgen05, tensor-bound), 'skinny' (HBM-bound by construction) or
    'tiny' (< 2 GFLOP: the 5x5 SPADE class-map convolutions, classifier heads, linears - launch-latency bound)."""
    return f"conv_{op}_" + ("tiny" if flops < 2e9 else spec.klass)


def conv_fwd_is_tma(spec: ConvSpec, n, h, w) -> bool:
    """True when dei2i_conv2d_fwd runs this geometry on the TMA / tcgen05 kernel (which can emit statistics / a padded output)."""
    if ACT_DTYPE == torch.float32:
        return False
    g = spec.geom(n, h, w)
    return bool(_lib.load().dei2i_conv2d_fwd_path(C.byref(g)))


def conv2d_fwd(x, w_fwd, spec: ConvSpec, bias=None, noise=None, noise_weight=None, residual=None, act=None,
               out_fp32=False, logical_hw=None, scale=None, alg_flops=None, want_stats=None, out_pad=0):
    """x: bf16 [N,Hs,Ws,Cin_ld]; logical_hw=(H,W) nearest-resizes x on the fly. Returns [N,Ho,Wo,Cout_ld].
    want_stats ('batch' | 'sample'): also returns the per-channel (mean, biased var) of the output over the batch / per image,
    emitted tile by tile from the convolution epilogue - or None where that kernel does not apply.  out_pad > 0: the output is
    [N,Ho+2p,Wo+2p,Cout_ld] with the reflect padding of its consumer in place (TMA kernel only; check conv_fwd_is_tma)."""
    _chk_act(x, "x")
    n, hs, ws, c = x.shape
    assert c == spec.cin_ld, (c, spec)
    h, w = logical_hw or (hs, ws)
    ho, wo = spec.out_hw(h, w)
    f32 = is_f32(x)
    y = torch.empty(n, ho + 2 * out_pad, wo + 2 * out_pad, spec.cout_ld, dtype=torch.float32 if (out_fp32 or f32) else torch.bfloat16, device=x.device)
    if residual is not None:
        _chk_act(residual, "residual")
        assert tuple(residual.shape) == (n, ho, wo, spec.cout_ld)
    if noise is not None:
        _chk_f32(noise, "noise")
        assert noise.numel() == n * ho * wo
    g = spec.geom(n, h, w, hs, ws)
    lib = _lib.load()
    if f32:  # FP32-parity path: one exact kernel for every geometry
        assert out_pad == 0 and w_fwd.dtype == torch.float32
        e = ConvEpilogue(_p(scale), _p(bias), _p(noise), _p(noise_weight), _p(residual), ACT[act], 1, None, 0)
        fl = alg_flops or spec.flops(n, ho, wo)
        with _Timed(_kind("fwd", spec, fl), fl, spec.tag(n, h, w)):
            check(lib.dei2i_conv2d_fwd_f32(_p(x), _p(w_fwd), _p(y), C.byref(g), C.byref(e), _stream()), "dei2i_conv2d_fwd_f32")
        _log_kernel("fwd", spec.tag(n, h, w))
        return y if want_stats is None else (y, None)
    partial = None
    if want_stats is not None and not out_fp32 and logical_hw is None and lib.dei2i_conv2d_fwd_path(C.byref(g)) \
            and (want_stats == "batch" or (ho * wo) % 128 == 0):
        partial = torch.empty((n * ho * wo + 127) // 128, 2, spec.cout_ld, dtype=torch.float32, device=x.device)
    e = ConvEpilogue(_p(scale), _p(bias), _p(noise), _p(noise_weight), _p(residual), ACT[act], int(out_fp32), _p(partial), int(out_pad))
    fl = alg_flops or spec.flops(n, ho, wo)
    with _Timed(_kind("fwd", spec, fl), fl, spec.tag(n, h, w)):
        check(lib.dei2i_conv2d_fwd(_p(x), _p(w_fwd), _p(y), C.byref(g), C.byref(e), _stream()), "dei2i_conv2d_fwd")
    _log_kernel("fwd", spec.tag(n, h, w))
    if want_stats is None:
        return y
    stats = None
    if partial is not None:
        groups = n if want_stats == "sample" else 1
        stats = stats_finalize(partial, (n * ho * wo) // groups, groups, spec.cout_ld)
    return y, stats


def conv2d_dgrad(dy, w_dgrad, spec: ConvSpec, in_hw, scale=None, alg_flops=None):
    """dy: bf16 [N,Ho,Wo,Cout_ld] -> dx bf16 [N,H,W,Cin_ld] (reflect padding folded back)."""
    _chk_act(dy, "dy")
    n = dy.shape[0]
    h, w = in_hw
    assert tuple(dy.shape[1:3]) == spec.out_hw(h, w) and dy.shape[3] == spec.cout_ld
    g = spec.geom(n, h, w)
    lib = _lib.load()
    flops = alg_flops or spec.flops(n, dy.shape[1], dy.shape[2])
    if is_f32(dy):
        p = spec.pad if spec.pad_mode == "reflect" else 0
        dxp = torch.empty(n, h + 2 * p, w + 2 * p, spec.cin_ld, dtype=torch.float32, device=dy.device)
        with _Timed(_kind("dgrad", spec, flops), flops, spec.tag(n, h, w)):
            check(lib.dei2i_conv2d_dgrad_f32(_p(dy), _p(w_dgrad), _p(dxp), C.byref(g), _p(scale), _stream()), "dei2i_conv2d_dgrad_f32")
        _log_kernel("dgrad", spec.tag(n, h, w))
        return reflect_fold(dxp, p) if p else dxp
    if spec.pad_mode == "reflect" and spec.pad > 0:
        p = spec.pad
        dxp = torch.empty(n, h + 2 * p, w + 2 * p, spec.cin_ld, dtype=torch.bfloat16, device=dy.device)
        with _Timed(_kind("dgrad", spec, flops), flops, spec.tag(n, h, w)):
            check(lib.dei2i_conv2d_dgrad(_p(dy), _p(w_dgrad), _p(dxp), C.byref(g), _p(scale), 0, _stream()), "dei2i_conv2d_dgrad")
        _log_kernel("dgrad", spec.tag(n, h, w))
        dx = torch.empty(n, h, w, spec.cin_ld, dtype=torch.bfloat16, device=dy.device)
        check(lib.dei2i_reflect_fold(_p(dxp), _p(dx), n, h, w, spec.cin_ld, p, _stream()), "dei2i_reflect_fold")
        return dx
    dx = torch.empty(n, h, w, spec.cin_ld, dtype=torch.bfloat16, device=dy.device)
    with _Timed(_kind("dgrad", spec, flops), flops, spec.tag(n, h, w)):
        check(lib.dei2i_conv2d_dgrad(_p(dy), _p(w_dgrad), _p(dx), C.byref(g), _p(scale), 0, _stream()), "dei2i_conv2d_dgrad")
    _log_kernel("dgrad", spec.tag(n, h, w))
    return dx


def reflect_pad(x, pad: int):
    """bf16 [N,H,W,C] -> [N,H+2p,W+2p,C], F.pad(mode='reflect')."""
    _chk_act(x, "x")
    n, h, w, c = x.shape
    xp = torch.empty(n, h + 2 * pad, w + 2 * pad, c, dtype=x.dtype, device=x.device)
    if is_f32(x):
        check(_lib.load().dei2i_reflect_pad_f32(_p(x), _p(xp), n, h, w, c, pad, _stream()), "dei2i_reflect_pad_f32")
        return xp
    with _Timed("stream_reflect_pad", x.numel() * 2 + xp.numel() * 2, ""):
        check(_lib.load().dei2i_reflect_pad(_p(x), _p(xp), n, h, w, c, pad, _stream()), "dei2i_reflect_pad")
    return xp


def reflect_fold(dxp, pad: int):
    """backward of reflect_pad: bf16 [N,H+2p,W+2p,C] -> [N,H,W,C]."""
    _chk_act(dxp, "dxp")
    n, hp, wp, c = dxp.shape
    h, w = hp - 2 * pad, wp - 2 * pad
    dx = torch.empty(n, h, w, c, dtype=dxp.dtype, device=dxp.device)
    if is_f32(dxp):
        check(_lib.load().dei2i_reflect_fold_f32(_p(dxp), _p(dx), n, h, w, c, pad, _stream()), "dei2i_reflect_fold_f32")
        return dx
    with _Timed("stream_reflect_fold", dxp.numel() * 2 + dx.numel() * 2, ""):
        check(_lib.load().dei2i_reflect_fold(_p(dxp), _p(dx), n, h, w, c, pad, _stream()), "dei2i_reflect_fold")
    return dx


_ws_cache = {}
_ws_retired = []  # outgrown workspaces stay alive: a captured CUDA graph may still hold their address


def _workspace(nbytes, device):
    key = (device.index, torch.cuda.current_stream().cuda_stream)
    buf = _ws_cache.get(key)
    if buf is None or buf.numel() < nbytes:
        if buf is not None:
            _ws_retired.append(buf)
        buf = torch.empty(max(nbytes, 64 << 20), dtype=torch.uint8, device=device)
        _ws_cache[key] = buf
    return buf


def conv2d_wgrad(x, dy, dw, spec: ConvSpec, accumulate=False, sn=None, logical_hw=None, alg_flops=None):
    """dw (fp32 [Cout,Cin,kh,kw]) = or += grad.  sn = (w_orig, u, v, inv_sigma) applies the spectral-norm backward."""
    _chk_act(x, "x")
    _chk_act(dy, "dy")
    _chk_f32(dw, "dw")
    n, hs, ws, _ = x.shape
    h, w = logical_hw or (hs, ws)
    g = spec.geom(n, h, w, hs, ws)
    lib = _lib.load()
    if is_f32(x):
        nbytes = lib.dei2i_conv2d_wgrad_f32_workspace(C.byref(g))
        wsb = _workspace(nbytes, x.device)
        w0, u, v, inv = sn if sn is not None else (None, None, None, None)
        fl = alg_flops or spec.flops(n, dy.shape[1], dy.shape[2])
        with _Timed(_kind("wgrad", spec, fl), fl, spec.tag(n, h, w)):
            check(lib.dei2i_conv2d_wgrad_f32(_p(x), _p(dy), _p(dw), spec.cin, 1.0 if accumulate else 0.0, C.byref(g), _p(w0), _p(u), _p(v),
                                             _p(inv), _p(wsb), wsb.numel(), _stream()), "dei2i_conv2d_wgrad_f32")
        _log_kernel("wgrad", spec.tag(n, h, w))
        return dw
    nbytes = lib.dei2i_conv2d_wgrad_workspace(C.byref(g), spec.cin)
    wsb = _workspace(nbytes, x.device)
    w0, u, v, inv = sn if sn is not None else (None, None, None, None)
    fl = alg_flops or spec.flops(n, dy.shape[1], dy.shape[2])
    with _Timed(_kind("wgrad", spec, fl), fl, spec.tag(n, h, w)):
        check(lib.dei2i_conv2d_wgrad(_p(x), _p(dy), _p(dw), spec.cin, 1.0 if accumulate else 0.0, C.byref(g), _p(w0), _p(u),
                                     _p(v), _p(inv), _p(wsb), wsb.numel(), _stream()), "dei2i_conv2d_wgrad")
    _log_kernel("wgrad", spec.tag(n, h, w))
    return dw


def spectral_norm(w, u, v, training: bool, eps: float = 1e-12):
    """In-place power iteration on (u, v) when training; returns the device tensor [1/sigma, sigma]."""
    rows = w.shape[0]
    cols = w.numel() // rows
    out = torch.empty(2, dtype=torch.float32, device=w.device)
    scratch = torch.empty(rows + 17 * cols, dtype=torch.float32, device=w.device)
    check(_lib.load().dei2i_spectral_norm(_p(w), _p(u), _p(v), rows, cols, int(training), eps, _p(out), _p(scratch),
                                          _stream()), "dei2i_spectral_norm")
    return out


def spectral_norm_multi(weights, training: bool, eps: float = 1e-12):
    """One power iteration (training) / sigma only (eval) for a list of (w, u, v) in four launches; u, v updated in place.
    Returns [(inv_sigma[2], u_snapshot, v_snapshot)] - the snapshots are what the spectral-norm backward needs once later
    forward passes have advanced u and v."""
    if not weights:
        return []
    dev = weights[0][0].device
    n = len(weights)
    rows = [w.shape[0] for w, _, _ in weights]
    cols = [w.numel() // w.shape[0] for w, _, _ in weights]
    out = torch.empty(n, 2, dtype=torch.float32, device=dev)
    al = lambda k: (k + 63) // 64 * 64  # 256-byte aligned snapshots: the wgrad second pass reads v with 16-byte loads
    snap = torch.empty(sum(al(r) for r in rows) + sum(al(c) for c in cols), dtype=torch.float32, device=dev)
    items = (_lib.SnItem * n)()
    off, res = 0, []
    for i, (w, u, v) in enumerate(weights):
        _chk_f32(w, "w")
        us = snap[off:off + rows[i]]
        off += al(rows[i])
        vs = snap[off:off + cols[i]]
        off += al(cols[i])
        it = items[i]
        it.w, it.u, it.v, it.inv_sigma, it.u_snap, it.v_snap = _p(w), _p(u), _p(v), _p(out[i]), _p(us), _p(vs)
        it.rows, it.cols = rows[i], cols[i]
        res.append((out[i], us, vs))
    lib = _lib.load()
    need = lib.dei2i_spectral_norm_multi_scratch(items, n)
    scratch = torch.empty(need, dtype=torch.float32, device=dev)
    check(lib.dei2i_spectral_norm_multi(items, n, int(training), eps, _p(scratch), need, _stream()), "dei2i_spectral_norm_multi")
    return res


def nchw_to_nhwc(x: torch.Tensor, c_ld: int = None):
    _chk_f32(x, "x")
    n, c, h, w = x.shape
    c_ld = c_ld or pad8(c)
    y = torch.empty(n, h, w, c_ld, dtype=ACT_DTYPE, device=x.device)
    if ACT_DTYPE == torch.float32:
        check(_lib.load().dei2i_nchw_to_nhwc_f32(_p(x), _p(y), n, c, h, w, c_ld, _stream()), "dei2i_nchw_to_nhwc_f32")
        return y
    check(_lib.load().dei2i_nchw_to_nhwc(_p(x), _p(y), n, c, h, w, c_ld, _stream()), "dei2i_nchw_to_nhwc")
    return y


def nhwc_to_nchw(x: torch.Tensor, c: int):
    _chk_act(x, "x")
    n, h, w, c_ld = x.shape
    y = torch.empty(n, c, h, w, dtype=torch.float32, device=x.device)
    if is_f32(x):
        check(_lib.load().dei2i_nhwc_to_nchw_f32(_p(x), _p(y), n, c, h, w, c_ld, _stream()), "dei2i_nhwc_to_nchw_f32")
        return y
    check(_lib.load().dei2i_nhwc_to_nchw(_p(x), _p(y), n, c, h, w, c_ld, _stream()), "dei2i_nhwc_to_nchw")
    return y


def rowpack_fwd(img: torch.Tensor, kw: int, pad: int):
    """fp32 NCHW [N,C<=8,H,W] -> bf16 [N,H+2*pad,W,64]: kw horizontally adjacent (reflect-padded) pixels per position."""
    _chk_f32(img, "img")
    n, c, h, w = img.shape
    xp = torch.empty(n, h + 2 * pad, w, 64, dtype=torch.bfloat16, device=img.device)
    with _Timed("stream_rowpack", img.numel() * 4 + xp.numel() * 2, ""):
        check(_lib.load().dei2i_rowpack_fwd(_p(img), _p(xp), n, c, h, w, kw, pad, _stream()), "dei2i_rowpack_fwd")
    return xp


def rowpack_bwd(dxp: torch.Tensor, c: int, kw: int, pad: int):
    """adjoint of rowpack_fwd: bf16 [N,H+2*pad,W,64] -> fp32 NCHW [N,c,H,W]."""
    _chk_act(dxp, "dxp")
    n, hp, w, _ = dxp.shape
    h = hp - 2 * pad
    dimg = torch.empty(n, c, h, w, dtype=torch.float32, device=dxp.device)
    with _Timed("stream_rowpack", dimg.numel() * 4 + dxp.numel() * 2, ""):
        check(_lib.load().dei2i_rowpack_bwd(_p(dxp), _p(dimg), n, c, h, w, kw, pad, _stream()), "dei2i_rowpack_bwd")
    return dimg


def upsample_pad(x, up: int = 1, pad=(0, 0, 0, 0)):
    """nearest up-sampling by ``up`` + zero frame pad = (top, left, bottom, right): bf16 [N,H,W,C] -> [N,up*H+pt+pb,up*W+pl+pr,C]."""
    _chk_act(x, "x")
    n, h, w, c = x.shape
    pt, pl, pb, pr = pad
    out = torch.empty(n, h * up + pt + pb, w * up + pl + pr, c, dtype=torch.bfloat16, device=x.device)
    check(_lib.load().dei2i_upsample_pad_fwd(_p(x), _p(out), n, h, w, c, up, pt, pl, pb, pr, _stream()), "dei2i_upsample_pad_fwd")
    return out


def upsample_pad_bwd(dout, in_hw, up: int = 1, pad=(0, 0, 0, 0)):
    _chk_act(dout, "dout")
    n, _, _, c = dout.shape
    h, w = in_hw
    pt, pl, pb, pr = pad
    assert tuple(dout.shape[1:3]) == (h * up + pt + pb, w * up + pl + pr)
    dx = torch.empty(n, h, w, c, dtype=torch.bfloat16, device=dout.device)
    check(_lib.load().dei2i_upsample_pad_bwd(_p(dout), _p(dx), n, h, w, c, up, pt, pl, pb, pr, _stream()), "dei2i_upsample_pad_bwd")
    return dx


def maxpool_fwd(x, k: int, stride: int, pad: int):
    """nn.MaxPool2d(k, stride, pad) -> (out, idx uint8 [N,Ho,Wo,C])."""
    _chk_act(x, "x")
    n, h, w, c = x.shape
    ho, wo = (h + 2 * pad - k) // stride + 1, (w + 2 * pad - k) // stride + 1
    out = torch.empty(n, ho, wo, c, dtype=torch.bfloat16, device=x.device)
    idx = torch.empty(n, ho, wo, c, dtype=torch.uint8, device=x.device)
    check(_lib.load().dei2i_maxpool_fwd(_p(x), _p(out), _p(idx), n, h, w, c, k, stride, pad, _stream()), "dei2i_maxpool_fwd")
    return out, idx


def maxpool_bwd(dout, idx, in_hw, k: int, stride: int, pad: int):
    _chk_act(dout, "dout")
    n, _, _, c = dout.shape
    h, w = in_hw
    dx = torch.empty(n, h, w, c, dtype=torch.bfloat16, device=dout.device)
    check(_lib.load().dei2i_maxpool_bwd(_p(dout), _p(idx), _p(dx), n, h, w, c, k, stride, pad, _stream()), "dei2i_maxpool_bwd")
    return dx


def avgpool_fwd(x, k: int, scale: float = 1.0):
    _chk_act(x, "x")
    n, h, w, c = x.shape
    out = torch.empty(n, h // k, w // k, c, dtype=torch.bfloat16, device=x.device)
    check(_lib.load().dei2i_avgpool_fwd(_p(x), _p(out), n, h, w, c, k, float(scale), _stream()), "dei2i_avgpool_fwd")
    return out


def avgpool_bwd(dout, k: int, scale: float = 1.0):
    _chk_act(dout, "dout")
    n, ho, wo, c = dout.shape
    dx = torch.empty(n, ho * k, wo * k, c, dtype=torch.bfloat16, device=dout.device)
    check(_lib.load().dei2i_avgpool_bwd(_p(dout), _p(dx), n, ho * k, wo * k, c, k, float(scale), _stream()), "dei2i_avgpool_bwd")
    return dx


def gap_fwd(x):
    """nn.AdaptiveAvgPool2d(1): bf16 [N,H,W,C] -> fp32 [N,C]."""
    _chk_act(x, "x")
    n, h, w, c = x.shape
    out = torch.empty(n, c, dtype=torch.float32, device=x.device)
    check(_lib.load().dei2i_gap_fwd(_p(x), _p(out), n, h * w, c, _stream()), "dei2i_gap_fwd")
    return out


def gap_bwd(dout, hw):
    _chk_f32(dout, "dout")
    n, c = dout.shape
    h, w = hw
    dx = torch.empty(n, h, w, c, dtype=torch.bfloat16, device=dout.device)
    check(_lib.load().dei2i_gap_bwd(_p(dout), _p(dx), n, h * w, c, _stream()), "dei2i_gap_bwd")
    return dx


def channel_stats(x, per_sample: bool):
    """-> fp32 [groups, 2, C]: per-channel (groups = 1) or per-(n, channel) mean and biased variance (deterministic, Welford-merged)."""
    _chk_act(x, "x")
    n, h, w, c = x.shape
    lib = _lib.load()
    stats = torch.empty(n if per_sample else 1, 2, c, dtype=torch.float32, device=x.device)
    if is_f32(x):
        check(lib.dei2i_channel_stats_f32(_p(x), n, h * w, c, int(per_sample), _p(stats), _stream()), "dei2i_channel_stats_f32")
        return stats
    need = lib.dei2i_channel_stats_scratch(n, h * w, c, int(per_sample))
    scratch = torch.empty(max(need, 1), dtype=torch.float32, device=x.device)
    with _Timed("stream_channel_stats", x.numel() * 2, ""):
        check(lib.dei2i_channel_stats(_p(x), n, h * w, c, int(per_sample), _p(stats), _p(scratch), need, _stream()), "dei2i_channel_stats")
    return stats


def stats_finalize(partial, rows_per_group: int, groups: int, c: int, rows_per_tile: int = 128):
    """Per-tile (mean, M2) pairs emitted by a convolution epilogue ([groups * T][2][ld]) -> fp32 [groups, 2, c] (mean, biased var)."""
    t_total, _, ld = partial.shape
    assert t_total % groups == 0, (partial.shape, groups)
    stats = torch.empty(groups, 2, c, dtype=torch.float32, device=partial.device)
    check(_lib.load().dei2i_stats_finalize(_p(partial), t_total // groups, rows_per_tile, rows_per_group, groups, c, ld, _p(stats),
                                           _stream()), "dei2i_stats_finalize")
    return stats


def bn_fwd(x, gamma, beta, running_mean, running_var, training, act=None, residual=None, momentum=0.1, eps=1e-5, out_pad=0,
           stats=None):
    """out_pad > 0: y is [N, H+2p, W+2p, C] and carries the reflect padding of its consumer (dei2i_bn_fwd_pad).
    stats: batch (mean, var) [1,2,C] already produced by the convolution epilogue (else computed here when training)."""
    _chk_act(x, "x")
    n, h, w, c = x.shape
    if training and stats is None:
        stats = channel_stats(x, False)
    if not training:
        stats = None
    save = torch.empty(2, c, dtype=torch.float32, device=x.device)
    if is_f32(x):
        y = torch.empty_like(x)
        check(_lib.load().dei2i_bn_fwd_f32(_p(x), _p(y), _p(residual), _p(stats), _p(gamma), _p(beta), _p(running_mean), _p(running_var),
                                           _p(save[0]), _p(save[1]), n, h * w, c, int(training), momentum, eps, ACT[act], _stream()),
              "dei2i_bn_fwd_f32")
        return (reflect_pad(y, out_pad) if out_pad else y), save
    if out_pad:
        y = torch.empty(n, h + 2 * out_pad, w + 2 * out_pad, c, dtype=x.dtype, device=x.device)
        with _Timed("stream_bn_fwd", x.numel() * 2 + y.numel() * 2 + (residual.numel() * 2 if residual is not None else 0), ""):
            check(_lib.load().dei2i_bn_fwd_pad(_p(x), _p(y), _p(residual), _p(stats), _p(gamma), _p(beta), _p(running_mean),
                                               _p(running_var), _p(save[0]), _p(save[1]), n, h, w, c, int(training), momentum, eps,
                                               ACT[act], int(out_pad), _stream()), "dei2i_bn_fwd_pad")
        return y, save
    y = torch.empty_like(x)
    with _Timed("stream_bn_fwd", x.numel() * 2 * (2 + int(residual is not None)), ""):
        check(_lib.load().dei2i_bn_fwd(_p(x), _p(y), _p(residual), _p(stats), _p(gamma), _p(beta), _p(running_mean),
                                       _p(running_var), _p(save[0]), _p(save[1]), n, h * w, c, int(training), momentum, eps,
                                       ACT[act], _stream()), "dei2i_bn_fwd")
    return y, save


def bn_bwd(x, dy, gamma, beta, save, training, act=None, dy_pad=0, fold_in_place=False):
    """-> (dx, dgamma, dbeta).  dy_pad > 0: dy is the gradient of the padded forward output, folded while it is read - or, with
    fold_in_place, its border is first folded into its own interior (dei2i_reflect_fold_inplace: 3 % of the pixels; dy is MODIFIED)
    and the kernels then read the interior through the padded pitch."""
    _chk_act(x, "x")
    _chk_act(dy, "dy")
    n, h, w, c = x.shape
    assert tuple(dy.shape) == (n, h + 2 * dy_pad, w + 2 * dy_pad, c), (dy.shape, x.shape, dy_pad)
    lib = _lib.load()
    if dy_pad and fold_in_place and not is_f32(x):
        check(lib.dei2i_reflect_fold_inplace(_p(dy), n, h, w, c, int(dy_pad), _stream()), "dei2i_reflect_fold_inplace")
        dy_pad = -int(dy_pad)
    dx = torch.empty_like(x)
    dbg = torch.empty(2, c, dtype=torch.float32, device=x.device)  # [dbeta | dgamma]
    if is_f32(x):
        if dy_pad:
            dy = reflect_fold(dy, abs(dy_pad))
        check(lib.dei2i_bn_bwd_f32(_p(x), _p(dy), _p(dx), _p(gamma), _p(beta), _p(save[0]), _p(save[1]), _p(dbg), n, h * w, c,
                                   int(training), ACT[act], _stream()), "dei2i_bn_bwd_f32")
        return dx, dbg[1], dbg[0]
    need = lib.dei2i_bn_bwd_scratch(n, h * w, c)
    scratch = torch.empty(need, dtype=torch.float32, device=x.device)
    with _Timed("stream_bn_bwd", x.numel() * 2 * 2 + dy.numel() * 2, ""):
        check(lib.dei2i_bn_bwd(_p(x), _p(dy), _p(dx), _p(gamma), _p(beta), _p(save[0]), _p(save[1]), _p(dbg[1]), _p(dbg[0]),
                               _p(scratch), need, n, h, w, c, int(training), ACT[act], int(dy_pad), _stream()), "dei2i_bn_bwd")
    return dx, dbg[1], dbg[0]


def spade_fwd(x, gb, up: int = 1, relu: bool = True, eps: float = 1e-5, stats=None):
    """x: [N,H/up,W/up,C]; gb: [N,H,W,2C] -> out [N,H,W,C] = relu(IN(x)*(1+gamma)+beta); also returns mean_rstd."""
    _chk_act(x, "x")
    _chk_act(gb, "gb")
    n, hx, wx, c = x.shape
    h, w = hx * up, wx * up
    assert tuple(gb.shape) == (n, h, w, 2 * c), (gb.shape, x.shape, up)
    if stats is None:
        stats = channel_stats(x, True)
    out = torch.empty(n, h, w, c, dtype=x.dtype, device=x.device)
    mr = torch.empty(n, 2, c, dtype=torch.float32, device=x.device)
    if is_f32(x):
        check(_lib.load().dei2i_spade_fwd_f32(_p(x), _p(gb), _p(out), _p(stats), _p(mr), n, h, w, c, up, int(relu), eps, _stream()),
              "dei2i_spade_fwd_f32")
        return out, mr
    with _Timed("stream_spade_fwd", x.numel() * 2 + gb.numel() * 2 + out.numel() * 2, ""):
        check(_lib.load().dei2i_spade_fwd(_p(x), _p(gb), _p(out), _p(stats), _p(mr), n, h, w, c, up, int(relu), eps, _stream()),
              "dei2i_spade_fwd")
    return out, mr


def spade_bwd(x, gb, out, dout, mean_rstd, up: int = 1, relu: bool = True):
    n, h, w, c = out.shape
    if is_f32(x):
        raise NotImplementedError("FP32-parity path: the dense SPADE backward (spatial label maps in training) is not built; "
                                  "training feeds 1x1 label maps (class-map path)")
    dgb = torch.empty_like(gb)
    dx = torch.empty_like(x)
    lib = _lib.load()
    need = lib.dei2i_image_reduce_scratch(n, h * w, c)
    scratch = torch.empty(need, dtype=torch.float32, device=x.device)
    with _Timed("stream_spade_bwd", x.numel() * 2 + 2 * gb.numel() * 2 + 2 * out.numel() * 2 + dx.numel() * 2, ""):
        check(lib.dei2i_spade_bwd(_p(x), _p(gb), _p(out), _p(dout), _p(dgb), _p(dx), _p(mean_rstd), _p(scratch), need, n, h, w,
                                  c, up, int(relu), _stream()), "dei2i_spade_bwd")
    return dgb, dx


def spade_cls_fwd(x, gb5, up: int = 1, relu: bool = True, eps: float = 1e-5, out_pad: int = 0, stats=None):
    """SPADE over a border-class map (constant label map): gb5 bf16 [N,5,5,2C]; otherwise like spade_fwd.  out_pad > 0: the
    output is [N,H+2p,W+2p,C] with the reflect padding of the consuming convolution already in place."""
    _chk_act(x, "x")
    _chk_act(gb5, "gb5")
    n, hx, wx, c = x.shape
    h, w = hx * up, wx * up
    assert tuple(gb5.shape) == (n, 5, 5, 2 * c), (gb5.shape, x.shape)
    if stats is None:
        stats = channel_stats(x, True)
    out = torch.empty(n, h + 2 * out_pad, w + 2 * out_pad, c, dtype=x.dtype, device=x.device)
    mr = torch.empty(n, 2, c, dtype=torch.float32, device=x.device)
    if is_f32(x):
        check(_lib.load().dei2i_spade_cls_fwd_f32(_p(x), _p(gb5), _p(out), _p(stats), _p(mr), n, h, w, c, up, int(relu), eps, out_pad,
                                                  _stream()), "dei2i_spade_cls_fwd_f32")
        return out, mr
    with _Timed("stream_spade_cls_fwd", x.numel() * 2 + out.numel() * 2, ""):
        check(_lib.load().dei2i_spade_cls_fwd(_p(x), _p(gb5), _p(out), _p(stats), _p(mr), n, h, w, c, up, int(relu), eps, out_pad,
                                              _stream()), "dei2i_spade_cls_fwd")
    return out, mr


def spade_cls_bwd(x, gb5, out, dout, mean_rstd, up: int = 1, relu: bool = True, out_pad: int = 0):
    """-> (dgb5 fp32 [N,5,5,2C], dx bf16).  out / dout carry ``out_pad`` pixels of reflect padding (folded on read)."""
    n, h, w, c = out.shape
    h, w = h - 2 * out_pad, w - 2 * out_pad
    assert dout.shape == out.shape
    dgb5 = torch.empty(n, 5, 5, 2 * c, dtype=torch.float32, device=x.device)
    dx = torch.empty_like(x)
    lib = _lib.load()
    if is_f32(x):
        red = torch.empty(n, 2, c, dtype=torch.float32, device=x.device)
        check(lib.dei2i_spade_cls_bwd_f32(_p(x), _p(gb5), _p(dout), _p(dgb5), _p(dx), _p(mean_rstd), _p(red), n, h, w, c, up, int(relu),
                                          out_pad, _stream()), "dei2i_spade_cls_bwd_f32")
        return dgb5, dx
    # (Walking the batch in L2-sized chunks, so that the second pass over x and dout hits L2, was measured and dropped: 48 MB chunks
    # 75.05 -> 79.1 ms per step, 80 MB 77.2 ms - the smaller grids and extra launches cost more than the HBM re-read.)
    need = lib.dei2i_spade_cls_bwd_scratch(n, h, w, c, up)
    scratch = torch.empty(need, dtype=torch.float32, device=x.device)
    with _Timed("stream_spade_cls_bwd", x.numel() * 2 * 2 + dout.numel() * 2, ""):
        check(lib.dei2i_spade_cls_bwd(_p(x), _p(gb5), _p(out), _p(dout), _p(dgb5), _p(dx), _p(mean_rstd), _p(scratch), need, n, h,
                                      w, c, up, int(relu), out_pad, _stream()), "dei2i_spade_cls_bwd")
    return dgb5, dx


def adain_act_fwd(x, scale, shift, slope: float = 0.2, eps: float = 1e-5, stats=None):
    """out = lrelu_slope(IN(x) * scale[n,c] + shift[n,c]); scale / shift fp32 [N,C].  Returns (out, mean_rstd)."""
    _chk_act(x, "x")
    _chk_f32(scale, "scale")
    _chk_f32(shift, "shift")
    n, h, w, c = x.shape
    assert tuple(scale.shape) == (n, c) and tuple(shift.shape) == (n, c), (scale.shape, shift.shape, x.shape)
    if stats is None:
        stats = channel_stats(x, True)
    out = torch.empty_like(x)
    mr = torch.empty(n, 2, c, dtype=torch.float32, device=x.device)
    check(_lib.load().dei2i_adain_act_fwd(_p(x), _p(out), _p(stats), _p(scale), _p(shift), _p(mr), n, h * w, c, float(slope), float(eps),
                                          _stream()), "dei2i_adain_act_fwd")
    return out, mr


def adain_act_bwd(x, out, dout, scale, mean_rstd, slope: float = 0.2):
    """-> (dx bf16, dscale fp32 [N,C], dshift fp32 [N,C])."""
    n, h, w, c = x.shape
    dx = torch.empty_like(x)
    red = torch.empty(n, 2, c, dtype=torch.float32, device=x.device)
    lib = _lib.load()
    need = lib.dei2i_image_reduce_scratch(n, h * w, c)
    scratch = torch.empty(need, dtype=torch.float32, device=x.device)
    check(lib.dei2i_adain_act_bwd(_p(x), _p(out), _p(dout), _p(dx), _p(scale), _p(mean_rstd), _p(red), _p(scratch), need, n, h * w, c,
                                  float(slope), _stream()), "dei2i_adain_act_bwd")
    return dx, red[:, 1], red[:, 0]


def act_fwd(x, act):
    _chk_act(x, "x")
    y = torch.empty_like(x)
    check(_lib.load().dei2i_act_fwd(_p(x), _p(y), x.numel(), ACT[act], _stream()), "dei2i_act_fwd")
    return y


def axpby(a, b, sa: float, sb: float = 0.0):
    """sa * a + sb * b over bf16 tensors of equal shape (b may be None)."""
    _chk_act(a, "a")
    if b is not None:
        _chk_act(b, "b")
        assert b.shape == a.shape, (a.shape, b.shape)
    out = torch.empty_like(a)
    check(_lib.load().dei2i_axpby(_p(a), _p(b), _p(out), a.numel(), float(sa), float(sb), _stream()), "dei2i_axpby")
    return out


def act_bwd(y, dy, act, want_dx=True, want_bias_grad=False, pad=0):
    """dx = dy * act'(y), bias grad = column sums of dx.  pad > 0: y and dy are consumer-padded [N,H+2p,W+2p,C] (conv2d_fwd with
    out_pad): dy is folded while it is read and dx comes back dense [N,H,W,C]."""
    _chk_act(dy, "dy")
    n, hp, wp, c = dy.shape
    h, w = hp - 2 * pad, wp - 2 * pad
    m = n * h * w
    dx = torch.empty(n, h, w, c, dtype=dy.dtype, device=dy.device) if want_dx else None
    bg = torch.empty(c, dtype=torch.float32, device=dy.device) if want_bias_grad else None
    if is_f32(dy):
        assert pad == 0
        check(_lib.load().dei2i_act_bwd_f32(_p(y), _p(dy), _p(dx), _p(bg), m, c, ACT[act], _stream()), "dei2i_act_bwd_f32")
        return dx, bg
    scratch = torch.empty(_lib.REDUCE_BLOCKS * c, dtype=torch.float32, device=dy.device) if want_bias_grad else None
    with _Timed("stream_act_bwd", dy.numel() * 2 * (int(y is not None) + 1 + int(want_dx)), ""):
        check(_lib.load().dei2i_act_bwd(_p(y), _p(dy), _p(dx), _p(bg), _p(scratch), m, c, ACT[act], h, w, int(pad), _stream()), "dei2i_act_bwd")
    return dx, bg


def noise_wgrad(dy, noise, dw, c_real, accumulate=False):
    c = dy.shape[3]
    m = dy.numel() // c
    if is_f32(dy):
        check(_lib.load().dei2i_noise_wgrad_f32(_p(dy), _p(noise), _p(dw), m, c, c_real, 1.0 if accumulate else 0.0, _stream()),
              "dei2i_noise_wgrad_f32")
        return dw
    scratch = torch.empty(_lib.REDUCE_BLOCKS, dtype=torch.float32, device=dy.device)
    with _Timed("stream_noise_wgrad", dy.numel() * 2, ""):
        check(_lib.load().dei2i_noise_wgrad(_p(dy), _p(noise), _p(dw), _p(scratch), m, c, c_real, 1.0 if accumulate else 0.0, _stream()),
              "dei2i_noise_wgrad")
    return dw


def compose_fwd(hd, x, cycle_gan=False, want_nhwc=True):
    """hd: fp32 [N,H,W,8] head logits; x: fp32 NCHW image -> (out NCHW fp32, prob [N,1,H,W] fp32, out NHWC bf16)."""
    n, h, w, _ = hd.shape
    out = torch.empty(n, 3, h, w, dtype=torch.float32, device=hd.device)
    prob = torch.empty(n, 1, h, w, dtype=torch.float32, device=hd.device)
    onhwc = torch.empty(n, h, w, 8, dtype=torch.bfloat16, device=hd.device) if want_nhwc else None
    with _Timed("stream_compose", hd.numel() * 4 + x.numel() * 4 + out.numel() * 4 + prob.numel() * 4, ""):
        check(_lib.load().dei2i_compose_fwd(_p(hd), _p(x), _p(out), _p(prob), _p(onhwc), n, h, w, int(cycle_gan), _stream()),
              "dei2i_compose_fwd")
    return out, prob, onhwc


def compose_bwd(hd, x, dout, dprob, cycle_gan=False, want_dx=True):
    n, h, w, _ = hd.shape
    dhd = torch.empty(n, h, w, 8, dtype=ACT_DTYPE, device=hd.device)
    dx = torch.empty(n, 3, h, w, dtype=torch.float32, device=hd.device) if want_dx else None
    if ACT_DTYPE == torch.float32:
        check(_lib.load().dei2i_compose_bwd_f32(_p(hd), _p(x), _p(dout), _p(dprob), _p(dhd), _p(dx), n, h, w, int(cycle_gan), _stream()),
              "dei2i_compose_bwd_f32")
        return dhd, dx
    with _Timed("stream_compose", hd.numel() * 4 + x.numel() * 4 + dhd.numel() * 2 + (dout.numel() * 4 if dout is not None else 0), ""):
        check(_lib.load().dei2i_compose_bwd(_p(hd), _p(x), _p(dout), _p(dprob), _p(dhd), _p(dx), n, h, w, int(cycle_gan),
                                            _stream()), "dei2i_compose_bwd")
    return dhd, dx


def bce_logits(logits, targets=None, target_const=0.0, want_grad=True, grad_scale=1.0):
    _chk_f32(logits, "logits")
    loss = torch.empty((), dtype=torch.float32, device=logits.device)
    grad = torch.empty_like(logits) if want_grad else None
    check(_lib.load().dei2i_bce_logits(_p(logits), _p(targets), float(target_const), logits.numel(), _p(loss), _p(grad),
                                       float(grad_scale), _stream()), "dei2i_bce_logits")
    return loss, grad


def l1(a, b=None, grad_a=None, grad_b=None, grad_scale=1.0, accumulate=False):
    _chk_f32(a, "a")
    loss = torch.empty((), dtype=torch.float32, device=a.device)
    scratch = torch.empty(_lib.REDUCE_BLOCKS, dtype=torch.float32, device=a.device)
    check(_lib.load().dei2i_l1(_p(a), _p(b), a.numel(), _p(loss), _p(grad_a), _p(grad_b), float(grad_scale), int(accumulate),
                               _p(scratch), _stream()), "dei2i_l1")
    return loss


def mse(a, b=None, grad_a=None, grad_b=None, grad_scale=1.0, accumulate=False):
    _chk_f32(a, "a")
    loss = torch.empty((), dtype=torch.float32, device=a.device)
    scratch = torch.empty(_lib.REDUCE_BLOCKS, dtype=torch.float32, device=a.device)
    check(_lib.load().dei2i_mse(_p(a), _p(b), a.numel(), _p(loss), _p(grad_a), _p(grad_b), float(grad_scale), int(accumulate),
                                _p(scratch), _stream()), "dei2i_mse")
    return loss


def ce_soft(logits, targets, want_grad=True, grad_scale=1.0):
    _chk_f32(logits, "logits")
    _chk_f32(targets, "targets")
    b, c = logits.shape
    loss = torch.empty((), dtype=torch.float32, device=logits.device)
    grad = torch.empty_like(logits) if want_grad else None
    check(_lib.load().dei2i_ce_soft(_p(logits), _p(targets), b, c, _p(loss), _p(grad), float(grad_scale), _stream()),
          "dei2i_ce_soft")
    return loss, grad


def adam(p, g, m, v, lr, beta1, beta2, eps, step):
    check(_lib.load().dei2i_adam(_p(p), _p(g), _p(m), _p(v), p.numel(), lr, beta1, beta2, eps, step, _stream()), "dei2i_adam")


def adam_dev(p, g, m, v, hyper, beta1, beta2, eps):
    """Adam / AdamW step with hyper = device fp32 [lr / (1 - beta1^t), 1 / sqrt(1 - beta2^t), 1 - lr * decoupled_weight_decay]
    (CUDA-graph friendly)."""
    check(_lib.load().dei2i_adam_dev(_p(p), _p(g), _p(m), _p(v), p.numel(), _p(hyper), beta1, beta2, eps, _stream()), "dei2i_adam_dev")


def rmsprop(p, g, sq, lr, alpha=0.99, eps=1e-8):
    check(_lib.load().dei2i_rmsprop(_p(p), _p(g), _p(sq), p.numel(), lr, alpha, eps, _stream()), "dei2i_rmsprop")


def rmsprop_dev(p, g, sq, hyper, alpha=0.99, eps=1e-8):
    """RMSprop step with the learning rate in device memory (hyper[0]) - CUDA-graph friendly."""
    check(_lib.load().dei2i_rmsprop_dev(_p(p), _p(g), _p(sq), p.numel(), _p(hyper), alpha, eps, _stream()), "dei2i_rmsprop_dev")


def clamp_(p, lo, hi):
    check(_lib.load().dei2i_clamp(_p(p), p.numel(), float(lo), float(hi), _stream()), "dei2i_clamp")


def ema_(p, p_ema, beta):
    check(_lib.load().dei2i_ema(_p(p), _p(p_ema), p.numel(), float(beta), _stream()), "dei2i_ema")


def _tdims(t):
    return (C.c_int * 4)(*[int(v) for v in t.shape]) if t is not None else None


def mask_token_fwd(img, coarse, h_shift, w_shift, patch, token=None, want_mask=True):
    """masked = img * m + token * (1 - m) (MaskToken.forward): img fp32 NCHW, coarse fp32 [N,gh,gw] Bernoulli grid, token None or a
    broadcastable fp32 4-D tensor.  Returns (masked, mask [N,1,H,W] or None)."""
    _chk_f32(img, "img")
    _chk_f32(coarse, "coarse")
    n, c, h, w = img.shape
    out = torch.empty_like(img)
    mask = torch.empty(n, 1, h, w, dtype=torch.float32, device=img.device) if want_mask else None
    if token is not None:
        _chk_f32(token, "token")
        assert token.dim() == 4
    check(_lib.load().dei2i_mask_token_fwd(_p(img), _p(coarse), _p(token), _tdims(token), _p(out), _p(mask), n, c, h, w, int(patch),
                                           int(h_shift), int(w_shift), coarse.shape[1], coarse.shape[2], _stream()), "dei2i_mask_token_fwd")
    return out, mask


def mask_mean(img, coarse, h_shift, w_shift, patch, mask_ratio):
    """'mean' mask token: [N,C,1,1] = mean_{h,w}(img * m) / mask_ratio (architecture.py:416-418)."""
    _chk_f32(img, "img")
    n, c, h, w = img.shape
    lib = _lib.load()
    tok = torch.empty(n, c, 1, 1, dtype=torch.float32, device=img.device)
    scratch = torch.empty(lib.dei2i_mask_token_scratch(n, c, h, w), dtype=torch.float32, device=img.device)
    check(lib.dei2i_mask_mean(_p(img), _p(coarse), _p(tok), _p(scratch), float(mask_ratio), n, c, h, w, int(patch), int(h_shift),
                              int(w_shift), coarse.shape[1], coarse.shape[2], _stream()), "dei2i_mask_mean")
    return tok


def mask_token_bwd(dy, coarse, h_shift, w_shift, patch, token_shape):
    """Gradient of a learnable mask token of shape ``token_shape`` ([1,1|C,1,1] or [1,1|C,H,W])."""
    _chk_f32(dy, "dy")
    n, c, h, w = dy.shape
    lib = _lib.load()
    dtok = torch.empty(tuple(token_shape), dtype=torch.float32, device=dy.device)
    scratch = torch.empty(lib.dei2i_mask_token_scratch(n, c, h, w), dtype=torch.float32, device=dy.device)
    check(lib.dei2i_mask_token_bwd(_p(dy), _p(coarse), (C.c_int * 4)(*[int(v) for v in token_shape]), _p(dtok), _p(scratch), n, c, h, w,
                                   int(patch), int(h_shift), int(w_shift), coarse.shape[1], coarse.shape[2], _stream()), "dei2i_mask_token_bwd")
    return dtok


def diffaug_fwd(x, params, bits: int):
    """DiffAugment forward on fp32 NCHW [N,C<=4,H,W]; params fp32 [N,8] (utils/diffaug.py), bits = dei2i_aug flags."""
    _chk_f32(x, "x")
    _chk_f32(params, "params")
    n, c, h, w = x.shape
    lib = _lib.load()
    y = torch.empty_like(x)
    scratch = torch.empty(lib.dei2i_diffaug_scratch(n, h, w), dtype=torch.float32, device=x.device)
    check(lib.dei2i_diffaug_fwd(_p(x), _p(y), _p(params), _p(scratch), n, c, h, w, int(bits), _stream()), "dei2i_diffaug_fwd")
    return y


def diffaug_bwd(dy, params, bits: int):
    _chk_f32(dy, "dy")
    n, c, h, w = dy.shape
    lib = _lib.load()
    dx = torch.empty_like(dy)
    scratch = torch.empty(lib.dei2i_diffaug_scratch(n, h, w), dtype=torch.float32, device=dy.device)
    check(lib.dei2i_diffaug_bwd(_p(dy), _p(dx), _p(params), _p(scratch), n, c, h, w, int(bits), _stream()), "dei2i_diffaug_bwd")
    return dx


def launch_count() -> int:
    return int(_lib.load().dei2i_launch_count())


def last_kernel() -> str:
    """Kernel variants the most recent conv2d_fwd / dgrad / wgrad call of this thread enqueued ('conv_tma<128,4,2>', ...)."""
    return _lib.load().dei2i_last_kernel().decode()


def set_tuning(key: str, value: int) -> int:
    return int(_lib.load().dei2i_set_tuning(key.encode(), int(value)))


def _apply_env_tuning():
    """DEI2I_TUNE="key=value,key=value": dei2i_set_tuning knobs from the environment (A/B runs of bench.py / the tools)."""
    spec = os.environ.get("DEI2I_TUNE", "")
    for kv in filter(None, (t.strip() for t in spec.split(","))):
        k, v = kv.split("=")
        set_tuning(k.strip(), int(v))


def ensure_env_tuning():
    global _env_tuned
    if not _env_tuned:
        _env_tuned = True
        _apply_env_tuning()
