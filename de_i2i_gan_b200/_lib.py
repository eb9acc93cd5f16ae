"""ctypes binding of libdei2i.so (the C ABI declared in include/dei2i.h).

There is no fallback: if the shared object is missing (not built) the import of any compute path raises.
"""
import ctypes as C
import os

REDUCE_BLOCKS = 2048  # DEI2I_REDUCE_BLOCKS
_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libdei2i.so")


class ConvGeom(C.Structure):
    _fields_ = [(n, C.c_int) for n in
                ("N", "H", "W", "Hs", "Ws", "Cin", "Cout", "Cout_ld", "kh", "kw", "stride", "pad", "pad_mode")]


class ConvEpilogue(C.Structure):
    _fields_ = [("scale", C.c_void_p), ("bias", C.c_void_p), ("noise", C.c_void_p), ("noise_weight", C.c_void_p), ("residual", C.c_void_p),
                ("act", C.c_int), ("out_fp32", C.c_int), ("stats_partial", C.c_void_p), ("out_pad", C.c_int)]


class SnItem(C.Structure):  # dei2i_sn_item
    _fields_ = [("w", C.c_void_p), ("u", C.c_void_p), ("v", C.c_void_p), ("inv_sigma", C.c_void_p), ("u_snap", C.c_void_p),
                ("v_snap", C.c_void_p), ("rows", C.c_int), ("cols", C.c_int)]


class PackItem(C.Structure):  # dei2i_pack_item
    _fields_ = [("w", C.c_void_p), ("w2", C.c_void_p), ("w_fwd", C.c_void_p), ("w_dgrad", C.c_void_p), ("rows_w", C.c_int),
                ("Cout", C.c_int), ("Cin_real", C.c_int), ("kh", C.c_int), ("kw", C.c_int), ("Cin", C.c_int), ("Cout_ld", C.c_int)]


_P, _I, _F, _LL, _SZ = C.c_void_p, C.c_int, C.c_float, C.c_longlong, C.c_size_t

# name -> (restype, argtypes); mirrors include/dei2i.h one to one
SIGNATURES = {
    "dei2i_version": (_I, []),
    "dei2i_last_error": (C.c_char_p, []),
    "dei2i_set_tuning": (_I, [C.c_char_p, _I]),
    "dei2i_launch_count": (_LL, []),
    "dei2i_last_kernel": (C.c_char_p, []),
    "dei2i_pack_weight": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _P, _P, _P]),
    "dei2i_conv2d_fwd": (_I, [_P, _P, _P, C.POINTER(ConvGeom), C.POINTER(ConvEpilogue), _P]),
    "dei2i_conv2d_fwd_path": (_I, [C.POINTER(ConvGeom)]),
    "dei2i_reflect_border": (_I, [_P, _I, _I, _I, _I, _I, _P]),
    "dei2i_conv2d_dgrad": (_I, [_P, _P, _P, C.POINTER(ConvGeom), _P, _I, _P]),
    "dei2i_reflect_fold": (_I, [_P, _P, _I, _I, _I, _I, _I, _P]),
    "dei2i_reflect_fold_inplace": (_I, [_P, _I, _I, _I, _I, _I, _P]),
    "dei2i_reflect_pad": (_I, [_P, _P, _I, _I, _I, _I, _I, _P]),
    "dei2i_pack_weight_multi": (_I, [C.POINTER(PackItem), _I, _P]),
    "dei2i_conv2d_wgrad_workspace": (_SZ, [C.POINTER(ConvGeom), _I]),
    "dei2i_conv2d_wgrad": (_I, [_P, _P, _P, _I, _F, C.POINTER(ConvGeom), _P, _P, _P, _P, _P, _SZ, _P]),
    "dei2i_spectral_norm": (_I, [_P, _P, _P, _I, _I, _I, _F, _P, _P, _P]),
    "dei2i_spectral_norm_multi_scratch": (_SZ, [C.POINTER(SnItem), _I]),
    "dei2i_spectral_norm_multi": (_I, [C.POINTER(SnItem), _I, _I, _F, _P, _SZ, _P]),
    "dei2i_nchw_to_nhwc": (_I, [_P, _P, _I, _I, _I, _I, _I, _P]),
    "dei2i_upsample_pad_fwd": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I, _P]),
    "dei2i_upsample_pad_bwd": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I, _P]),
    "dei2i_maxpool_fwd": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _P]),
    "dei2i_maxpool_bwd": (_I, [_P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _P]),
    "dei2i_avgpool_fwd": (_I, [_P, _P, _I, _I, _I, _I, _I, _F, _P]),
    "dei2i_avgpool_bwd": (_I, [_P, _P, _I, _I, _I, _I, _I, _F, _P]),
    "dei2i_gap_fwd": (_I, [_P, _P, _I, _I, _I, _P]),
    "dei2i_gap_bwd": (_I, [_P, _P, _I, _I, _I, _P]),
    "dei2i_adain_act_fwd": (_I, [_P, _P, _P, _P, _P, _P, _I, _I, _I, _F, _F, _P]),
    "dei2i_adain_act_bwd": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _SZ, _I, _I, _I, _F, _P]),
    "dei2i_act_fwd": (_I, [_P, _P, _LL, _I, _P]),
    "dei2i_axpby": (_I, [_P, _P, _P, _LL, _F, _F, _P]),
    "dei2i_rowpack_fwd": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _P]),
    "dei2i_rowpack_bwd": (_I, [_P, _P, _I, _I, _I, _I, _I, _I, _P]),
    "dei2i_nhwc_to_nchw": (_I, [_P, _P, _I, _I, _I, _I, _I, _P]),
    "dei2i_channel_stats_scratch": (_SZ, [_I, _I, _I, _I]),
    "dei2i_channel_stats": (_I, [_P, _I, _I, _I, _I, _P, _P, _SZ, _P]),
    "dei2i_channel_stats_static_max_floats": (_SZ, []),
    "dei2i_channel_stats_static": (_I, [_P, _I, _I, _I, _I, _P, _P]),
    "dei2i_stats_finalize": (_I, [_P, _I, _I, _LL, _I, _I, _I, _P, _P]),
    "dei2i_bn_fwd": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _F, _F, _I, _P]),
    "dei2i_bn_fwd_pad": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _F, _F, _I, _I, _P]),
    "dei2i_bn_bwd_scratch": (_SZ, [_I, _I, _I]),
    "dei2i_bn_bwd": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _SZ, _I, _I, _I, _I, _I, _I, _I, _P]),
    "dei2i_spade_fwd": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _F, _P]),
    "dei2i_image_reduce_scratch": (_SZ, [_I, _I, _I]),
    "dei2i_spade_bwd": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _SZ, _I, _I, _I, _I, _I, _I, _P]),
    "dei2i_spade_cls_fwd": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _F, _I, _P]),
    "dei2i_spade_cls_bwd_scratch": (_SZ, [_I, _I, _I, _I, _I]),
    "dei2i_spade_cls_bwd": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _SZ, _I, _I, _I, _I, _I, _I, _I, _P]),
    "dei2i_act_bwd": (_I, [_P, _P, _P, _P, _P, _LL, _I, _I, _I, _I, _I, _P]),
    "dei2i_noise_wgrad": (_I, [_P, _P, _P, _P, _LL, _I, _I, _F, _P]),
    "dei2i_compose_fwd": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "dei2i_compose_bwd": (_I, [_P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "dei2i_bce_logits": (_I, [_P, _P, _F, _LL, _P, _P, _F, _P]),
    "dei2i_l1": (_I, [_P, _P, _LL, _P, _P, _P, _F, _I, _P, _P]),
    "dei2i_mse": (_I, [_P, _P, _LL, _P, _P, _P, _F, _I, _P, _P]),
    "dei2i_ce_soft": (_I, [_P, _P, _I, _I, _P, _P, _F, _P]),
    "dei2i_adam": (_I, [_P, _P, _P, _P, _LL, _F, _F, _F, _F, _I, _P]),
    "dei2i_adam_dev": (_I, [_P, _P, _P, _P, _LL, _P, _F, _F, _F, _P]),
    "dei2i_rmsprop": (_I, [_P, _P, _P, _LL, _F, _F, _F, _P]),
    "dei2i_rmsprop_dev": (_I, [_P, _P, _P, _LL, _P, _F, _F, _P]),
    "dei2i_clamp": (_I, [_P, _LL, _F, _F, _P]),
    "dei2i_ema": (_I, [_P, _P, _LL, _F, _P]),
    "dei2i_pack_weight_f32": (_I, [_P, _I, _I, _I, _I, _I, _I, _P, _P, _P]),
    "dei2i_conv2d_fwd_f32": (_I, [_P, _P, _P, C.POINTER(ConvGeom), C.POINTER(ConvEpilogue), _P]),
    "dei2i_conv2d_dgrad_f32": (_I, [_P, _P, _P, C.POINTER(ConvGeom), _P, _P]),
    "dei2i_conv2d_wgrad_f32_workspace": (_SZ, [C.POINTER(ConvGeom)]),
    "dei2i_conv2d_wgrad_f32": (_I, [_P, _P, _P, _I, _F, C.POINTER(ConvGeom), _P, _P, _P, _P, _P, _SZ, _P]),
    "dei2i_nchw_to_nhwc_f32": (_I, [_P, _P, _I, _I, _I, _I, _I, _P]),
    "dei2i_nhwc_to_nchw_f32": (_I, [_P, _P, _I, _I, _I, _I, _I, _P]),
    "dei2i_channel_stats_f32": (_I, [_P, _I, _I, _I, _I, _P, _P]),
    "dei2i_bn_fwd_f32": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _F, _F, _I, _P]),
    "dei2i_bn_bwd_f32": (_I, [_P, _P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "dei2i_spade_cls_fwd_f32": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _F, _I, _P]),
    "dei2i_spade_cls_bwd_f32": (_I, [_P, _P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _I, _P]),
    "dei2i_spade_fwd_f32": (_I, [_P, _P, _P, _P, _P, _I, _I, _I, _I, _I, _I, _F, _P]),
    "dei2i_reflect_pad_f32": (_I, [_P, _P, _I, _I, _I, _I, _I, _P]),
    "dei2i_reflect_fold_f32": (_I, [_P, _P, _I, _I, _I, _I, _I, _P]),
    "dei2i_act_bwd_f32": (_I, [_P, _P, _P, _P, _LL, _I, _I, _P]),
    "dei2i_noise_wgrad_f32": (_I, [_P, _P, _P, _LL, _I, _I, _F, _P]),
    "dei2i_compose_bwd_f32": (_I, [_P, _P, _P, _P, _P, _P, _I, _I, _I, _I, _P]),
    "dei2i_mask_token_scratch": (_SZ, [_I, _I, _I, _I]),
    "dei2i_mask_token_fwd": (_I, [_P, _P, _P, C.POINTER(C.c_int), _P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I, _P]),
    "dei2i_mask_mean": (_I, [_P, _P, _P, _P, _F, _I, _I, _I, _I, _I, _I, _I, _I, _I, _P]),
    "dei2i_mask_token_bwd": (_I, [_P, _P, C.POINTER(C.c_int), _P, _P, _I, _I, _I, _I, _I, _I, _I, _I, _I, _P]),
    "dei2i_augment_scratch": (_SZ, [_I, _I]),
    "dei2i_augment_train": (_I, [_P, _P, _P, _P, _P, _P, _I, _I, _I, _P]),
    "dei2i_diffaug_scratch": (_SZ, [_I, _I, _I]),
    "dei2i_diffaug_fwd": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
    "dei2i_diffaug_bwd": (_I, [_P, _P, _P, _P, _I, _I, _I, _I, _I, _P]),
}

_lib = None


def load():
    """Loads libdei2i.so and declares every prototype. Raises if the library was not built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(there is no CPU / PyTorch fallback for the CUDA path)")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the .so does not export a declared symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int, what: str):
    if rc != 0:
        msg = load().dei2i_last_error().decode("utf-8", "replace")
        raise RuntimeError(f"{what} failed (status {rc}): {msg}")
