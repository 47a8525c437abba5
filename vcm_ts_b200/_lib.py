"""ctypes binding of libvcmts_b200.so (the C-ABI in include/vcmts_b200.h).

There is no fallback: if the library is missing it is built (nvcc), and if that fails, or a kernel
is requested without a B200 visible, the call raises.
"""
import ctypes
import os
from ctypes import (POINTER, Structure, c_char_p, c_float, c_int, c_int16, c_int32, c_int64,
                    c_uint8, c_void_p)

HERE = os.path.dirname(os.path.abspath(__file__))
SO = os.environ.get("VCM_B200_SO") or os.path.join(HERE, "libvcmts_b200.so")   # override: A/B runs of two builds

F32, BF16 = 0, 1
CONV_PRE_ACT, CONV_POST_ACT, CONV_PIXEL_SHUFFLE = 1, 2, 4
CONV_RES, CONV_RES_SCALE, CONV_RES2, CONV_OUT2, CONV_SPLIT3, CONV_PDL, CONV_SE_SUM = 8, 16, 32, 64, 128, 256, 512


class ConvDesc(Structure):
    _fields_ = [("N", c_int32), ("H", c_int32), ("W", c_int32), ("Cin", c_int32), ("Cout", c_int32),
                ("K", c_int32), ("stride", c_int32), ("in_cstride", c_int32),
                ("out_cstride", c_int32), ("res_cstride", c_int32), ("res2_cstride", c_int32),
                ("out2_cstride", c_int32),
                ("flags", c_int32), ("pre_slope", c_float), ("post_slope", c_float),
                ("out2_slope", c_float),
                ("in_dtype", c_int32), ("out_dtype", c_int32), ("in_pstride", c_int32)]


class DpDesc(Structure):
    _fields_ = [("N", c_int32), ("h", c_int32), ("w", c_int32), ("C", c_int32),
                ("y_cstride", c_int32), ("fusion_cstride", c_int32), ("prior_cstride", c_int32),
                ("params_cstride", c_int32), ("yhat_cstride", c_int32), ("dtype", c_int32),
                ("mode", c_int32), ("log_scale_min", c_float), ("log_scale_step", c_float)]


_P = c_void_p
_PROTOS = {
    "vcm_last_error": (c_char_p, []),
    "vcm_version": (c_int, []),
    "vcm_device_ok": (c_int, []),
    "vcm_pack_conv_weight": (c_int, [_P, _P, c_int32, c_int32, c_int32, c_int32, _P, _P]),
    "vcm_conv2d_fwd": (c_int, [POINTER(ConvDesc), _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "vcm_conv2d_tc_supported": (c_int, [POINTER(ConvDesc)]),
    "vcm_conv2d_tc_weight_bytes": (c_int64, [c_int32, c_int32, c_int32]),
    "vcm_pack_conv_weight_umma": (c_int, [_P, _P, c_int32, c_int32, c_int32, c_int32, _P, _P]),
    "vcm_conv2d_tc_fwd": (c_int, [POINTER(ConvDesc), _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "vcm_split_bf16x3": (c_int, [_P, _P, c_int64, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32,
                                 c_float, _P]),
    "vcm_pack_conv_weight_umma_split3": (c_int, [_P, _P, c_int32, c_int32, c_int32, c_int32, _P, _P]),
    "vcm_nchw_to_nhwc": (c_int, [_P, _P, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32, _P]),
    "vcm_u8hwc_to_frame": (c_int, [_P, _P, c_int32, c_int32, c_int32, c_int32, c_int32, _P]),
    "vcm_frame_to_u8hwc": (c_int, [_P, _P, c_int32, c_int32, c_int32, c_int32, c_int32, _P]),
    "vcm_enh_residual": (c_int, [_P, _P, _P, c_int32, c_int32, c_int32, c_int32, _P, _P, c_int32, _P]),
    "vcm_enh_fuse": (c_int, [_P, _P, _P, c_int32, c_int32, c_int32, c_int32, _P, _P, _P, _P, c_int32, _P]),
    "vcm_nhwc_to_nchw": (c_int, [_P, _P, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32,
                                 c_int32, c_float, c_float, _P]),
    "vcm_copy_channels": (c_int, [_P, _P, c_int64, c_int32, c_int32, c_int32, c_int32, c_int32, _P]),
    "vcm_scale_channels": (c_int, [_P, _P, _P, c_int32, c_int64, c_int32, c_int32, c_int32, c_int32,
                                   c_int32, _P]),
    "vcm_maxpool2": (c_int, [_P, _P, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32, _P]),
    "vcm_se_partial_count": (c_int32, [c_int64]),
    "vcm_se_scale": (c_int, [_P, _P, _P, _P, _P, c_int32, c_int64, c_int32, c_int32, c_int32, c_int32, _P]),
    "vcm_conv2d_se_sum_rows": (c_int32, [POINTER(ConvDesc)]),
    "vcm_se_scale_from_sums": (c_int, [_P, c_int32, _P, _P, _P, c_int64, c_int32, c_int32, _P]),
    "vcm_avgpool2": (c_int, [_P, _P, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32, _P]),
    "vcm_flow_up2": (c_int, [_P, _P, c_int32, c_int32, c_int32, c_int32, c_int32, c_float, _P]),
    "vcm_flow_down2": (c_int, [_P, _P, c_int32, c_int32, c_int32, c_int32, c_int32, c_float, _P]),
    "vcm_spynet_pyramid": (c_int, [_P, _P, _P, _P, c_int32, c_int32, c_int32, _P]),
    "vcm_spynet_level_input": (c_int, [_P, _P, _P, _P, c_int32, c_int32, c_int32, c_int32, _P]),
    "vcm_flow_warp": (c_int, [_P, _P, _P, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32,
                              c_int32, c_int32, _P]),
    "vcm_conv2d_wgrad_scratch_floats": (c_int64, [POINTER(ConvDesc)]),
    "vcm_conv2d_wgrad": (c_int, [POINTER(ConvDesc), _P, _P, _P, _P, _P]),
    "vcm_flow_warp_bwd": (c_int, [_P, _P, _P, _P, _P, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32,
                                  c_int32, c_int32, _P]),
    "vcm_dual_prior_pass0": (c_int, [POINTER(DpDesc), _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "vcm_dual_prior_pass1": (c_int, [POINTER(DpDesc), _P, _P, _P, _P, _P, _P, _P, _P, _P, _P, _P]),
    "vcm_round_latent": (c_int, [_P, _P, _P, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32,
                                 c_int32, _P]),
    "vcm_symbols_to_latent": (c_int, [_P, _P, c_int32, c_int32, c_int32, c_int32, c_int32, c_int32, _P]),
    "vcm_bits_partial_count": (c_int32, [c_int64]),
    "vcm_latent_bits": (c_int, [_P, _P, _P, _P, c_int32, c_int64, c_int32, _P]),
    "vcm_factorized_bits": (c_int, [_P, _P, _P, _P, c_int32, c_int64, c_int32, c_int32, c_int32, _P]),
    "vcm_sq_err_sum": (c_int, [_P, _P, _P, _P, c_int32, c_int64, _P]),
    "vcm_sq_err_sum_noalloc": (c_int, [_P, _P, _P, c_int32, c_int64, _P]),
    "vcm_cdf_tables_new": (_P, [_P, _P, _P, c_int32, c_int32]),
    "vcm_cdf_tables_free": (None, [_P]),
    "vcm_rans_enc_new": (_P, []),
    "vcm_rans_enc_free": (None, [_P]),
    "vcm_rans_enc_reset": (None, [_P]),
    "vcm_rans_enc_push": (c_int, [_P, _P, _P, c_int64, _P]),
    "vcm_rans_enc_push_channels": (c_int, [_P, _P, c_int32, c_int64, _P]),
    "vcm_rans_enc_flush": (c_int64, [_P, POINTER(c_void_p)]),
    "vcm_rans_dec_new": (_P, []),
    "vcm_rans_dec_free": (None, [_P]),
    "vcm_rans_dec_set_stream": (c_int, [_P, c_char_p, c_int64]),
    "vcm_rans_dec_decode": (c_int, [_P, _P, c_int64, _P, _P]),
    "vcm_rans_dec_decode_channels": (c_int, [_P, c_int32, c_int64, _P, _P]),
    "vcm_pmf_to_quantized_cdf": (c_int, [_P, c_int32, c_int32, _P]),
}

EXPORTS = tuple(_PROTOS)
_lib = None


def lib():
    """Load (building first if needed) libvcmts_b200.so.  Raises if it cannot be had."""
    global _lib
    if _lib is None:
        if not os.path.exists(SO):
            from . import build as _build
            _build.build()
        L = ctypes.CDLL(SO)
        for name, (res, args) in _PROTOS.items():
            fn = getattr(L, name)          # AttributeError if the .so lacks a declared symbol
            fn.restype = res
            fn.argtypes = args
        _lib = L
    return _lib


class VcmError(RuntimeError):
    pass


def check(rc, what=""):
    if rc != 0:
        msg = lib().vcm_last_error()
        raise VcmError(f"{what} failed (rc={rc}): {msg.decode() if msg else ''}")
