"""ctypes binding of libb2srgan.so (the C ABI declared in include/b2srgan.h).

There is no fallback: if the CUDA extension is missing or a call fails, this raises.
"""
import ctypes
import os

from . import _build

_lib = None

ACT_NONE, ACT_RELU, ACT_LEAKY, ACT_PRELU, ACT_TANH = 0, 1, 2, 3, 4


class B2Error(RuntimeError):
    pass


class Epilogue(ctypes.Structure):
    _fields_ = [
        ("scale", ctypes.c_void_p),
        ("shift", ctypes.c_void_p),
        ("alpha", ctypes.c_void_p),
        ("act", ctypes.c_int),
        ("leaky", ctypes.c_float),
        ("residual", ctypes.c_void_p),
        ("out", ctypes.c_void_p),
        ("out_fp32", ctypes.c_int),
        ("f16", ctypes.c_int),
        ("gate", ctypes.c_void_p),
        ("gate_slope", ctypes.c_float),
        ("stats_part", ctypes.c_void_p),
    ]


_P, _I, _L, _F = ctypes.c_void_p, ctypes.c_int, ctypes.c_int64, ctypes.c_float

_SIGNATURES = {
    "b2_last_error": (ctypes.c_char_p, []),
    "b2_version": (ctypes.c_int, []),
    "b2_sizeof_epilogue": (ctypes.c_int, []),
    "b2_epilogue_field_count": (ctypes.c_int, []),
    "b2_epilogue_field_offset": (ctypes.c_int, [ctypes.c_int]),
    "b2_set_pair_kernels": (ctypes.c_int, [ctypes.c_int]),
    "b2_set_pdl": (ctypes.c_int, [ctypes.c_int]),
    "b2_conv2d_fprop": (ctypes.c_int, [ctypes.c_void_p] + [ctypes.c_int] * 4 + [ctypes.c_void_p] + [ctypes.c_int] * 4 +
                        [ctypes.POINTER(Epilogue), ctypes.c_void_p]),
    "b2_conv2d_transpose_s2_fprop": (ctypes.c_int, [ctypes.c_void_p] + [ctypes.c_int] * 4 + [ctypes.c_void_p] +
                                     [ctypes.c_int] * 3 + [ctypes.POINTER(Epilogue), ctypes.c_void_p]),
    "b2_conv2d_rows_fprop": (ctypes.c_int, [ctypes.c_void_p] + [ctypes.c_int] * 4 + [ctypes.c_void_p] + [ctypes.c_int] * 3 +
                             [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]),
    "b2_conv2d_scatter_fprop": (ctypes.c_int, [ctypes.c_void_p] + [ctypes.c_int] * 4 + [ctypes.c_void_p] + [ctypes.c_int] * 3 +
                                [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_void_p, ctypes.c_void_p]),
    "b2_conv2d_direct": (ctypes.c_int, [ctypes.c_void_p] + [ctypes.c_int] * 4 + [ctypes.c_void_p] + [ctypes.c_int] * 5 +
                         [ctypes.POINTER(Epilogue), ctypes.c_void_p]),
    "b2_fold_taps": (ctypes.c_int, [ctypes.c_void_p] + [ctypes.c_int] * 9 + [ctypes.POINTER(ctypes.c_int)] * 2 +
                     [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int, ctypes.c_void_p]),
    "b2_gather_cast": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p, ctypes.c_int,
                                      ctypes.c_void_p]),
    "b2_resize_bilinear_into": (ctypes.c_int, [ctypes.c_void_p] + [ctypes.c_int] * 5 + [ctypes.c_void_p] + [ctypes.c_int] * 3 + [ctypes.c_void_p]),
    "b2_u8_to_f32_norm": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]),
    "b2_count_saturated": (ctypes.c_int, [_P, _L, _I, _P, _P]),
    "b2_f32_to_u8_denorm": (ctypes.c_int, [ctypes.c_void_p, ctypes.c_void_p, ctypes.c_int64, ctypes.c_void_p]),
    "b2_gate_mul": (ctypes.c_int, [_P, _I, _I, _P, _P, _L, _I, _I, _P]),
    "b2_resize_nearest_bilinear_pair": (ctypes.c_int, [_P, _I, _I, _I, _I, _I, _P, _P]),
    "b2_atanh_conv2d_transpose_small": (ctypes.c_int, [_P, _I, _I, _I, _I, _P, _P, _I, _I, _P, _I, _P]),
    # ---- training step
    "b2_conv2d_s2_dgrad": (ctypes.c_int, [_P] + [_I] * 4 + [_P] + [_I] * 5 + [ctypes.POINTER(Epilogue), _P]),
    "b2_conv2d_wgrad": (ctypes.c_int, [_P] + [_I] * 4 + [_P] + [_I] * 4 + [_P, _I, _P]),
    "b2_conv2d_fprop_pad": (ctypes.c_int, [_P] + [_I] * 4 + [_P] + [_I] * 8 + [ctypes.POINTER(Epilogue), _P]),
    "b2_conv2d_strided_dgrad": (ctypes.c_int, [_P] + [_I] * 4 + [_P] + [_I] * 8 + [ctypes.POINTER(Epilogue), _P]),
    "b2_conv2d_wgrad_pad": (ctypes.c_int, [_P] + [_I] * 4 + [_P] + [_I] * 8 + [_P, _I, _P]),
    "b2_colstats": (ctypes.c_int, [_P, _L, _I, _I, _P, _P, _P]),
    "b2_bn_finalize": (ctypes.c_int, [_P, _P, _F, _P, _P, _F, _F, _P, _P, _P, _P, _P, _P, _I, _P]),
    "b2_bn_fold": (ctypes.c_int, [_P, _P, _P, _P, _P, _F, _P, _P, _I, _P]),
    "b2_affine_act": (ctypes.c_int, [_P, _L, _I, _I, _P, _P, _P, _I, _F, _P, _P, _P]),
    "b2_bn_bwd_reduce": (ctypes.c_int, [_P, _P, _L, _I, _I, _P, _P, _P, _I, _F, _P, _P, _P, _P, _P, _P]),
    "b2_bn_bwd_apply": (ctypes.c_int, [_P, _P, _L, _I, _I, _P, _P, _P, _I, _F, _P, _P, _P, _P, _F, _P, _P]),
    "b2_sm_count": (ctypes.c_int, []),
    "b2_bn_part_rows": (ctypes.c_int, []),
    "b2_bn_stats_partial": (ctypes.c_int, [_P, _L, _I, _I, _P, ctypes.POINTER(ctypes.c_int), _P]),
    "b2_bn_fwd_finalize": (ctypes.c_int, [_P, _I, _I, _F, ctypes.POINTER(ctypes.c_void_p), _I, _I, ctypes.c_uint, _P, _P, _P, _P, _F, _F,
                                          _P, _P, _P, _P, _F, _P]),
    "b2_bn_fwd_stats_static": (ctypes.c_int, [_P, _L, _I, _I, _F, _P, _P, _F, _P, _P, _P]),
    "b2_bn_bwd_partial": (ctypes.c_int, [_P, _P, _L, _I, _I, _P, _P, _I, _F, _P, ctypes.POINTER(ctypes.c_int), _P]),
    "b2_bn_bwd_finalize": (ctypes.c_int, [_P, _I, _I, ctypes.POINTER(ctypes.c_void_p), _I, _I, ctypes.c_uint, _P, _P, _P, _P, _P, _P,
                                          _F, _P]),
    "b2_maxpool2x2": (ctypes.c_int, [_P, _I, _I, _I, _I, _I, _P, _P]),
    "b2_maxpool2x2_bwd": (ctypes.c_int, [_P, _P, _P, _I, _I, _I, _I, _I, _I, _P, _P]),
    "b2_loss_pair": (ctypes.c_int, [_P, _P, _L, _I, _I, _F, _F, _P, _P, _P]),
    "b2_mean_prod": (ctypes.c_int, [_P, _P, _F, _L, _F, _F, _P, _P, _P]),
    "b2_score_act": (ctypes.c_int, [_P, _P, _P, _L, _I, _P]),
    "b2_peer_max_elems": (ctypes.c_int, []),
    "b2_peer_alloc": (ctypes.c_int, [ctypes.POINTER(ctypes.c_void_p), ctypes.c_char_p]),
    "b2_peer_open": (ctypes.c_int, [ctypes.c_char_p, ctypes.POINTER(ctypes.c_void_p)]),
    "b2_peer_close": (ctypes.c_int, [_P, _I]),
    "b2_peer_exchange_probe": (ctypes.c_int, [_P, _I, ctypes.POINTER(ctypes.c_void_p), _I, _I, ctypes.c_uint, _I, _P, _P]),
    "b2_peer_allreduce_f32": (ctypes.c_int, [_P, _I, ctypes.POINTER(ctypes.c_void_p), _I, _I, ctypes.c_uint, _P, _P]),
    "b2_tanh_bwd": (ctypes.c_int, [_P, _P, _P, _P, _L, _P, _P]),
    "b2_gather_f32": (ctypes.c_int, [_P, _P, _L, _P, _I, _P]),
    "b2_colsum_f32": (ctypes.c_int, [_P, _L, _I, _P, _P]),
    "b2_dense_fwd": (ctypes.c_int, [_P, _I, _I, _I, _I, _P, _P, _P, _I, _P]),
    "b2_dense_bwd_data": (ctypes.c_int, [_P, _I, _I, _I, _I, _P, _P, _I, _P]),
    "b2_dense_bwd_weight": (ctypes.c_int, [_P, _I, _P, _I, _I, _I, _I, _P, _P, _P]),
    "b2_cast_scale": (ctypes.c_int, [_P, _I, _L, _F, _P, _I, _P]),
    "b2_adam": (ctypes.c_int, [_P, _P, _P, _P, _L, _F, _F, _F, _F, _F, _P]),
    "b2_adam_guarded": (ctypes.c_int, [_P, _P, _P, _P, _L, _F, _F, _F, _F, _F, _P, _P, _I, _P]),
}


def exported_symbols():
    """Names include/b2srgan.h declares; tests check the loaded library exports each of them."""
    return sorted(_SIGNATURES)


def lib_path():
    return _build.LIB_PATH


def load():
    """Loads libb2srgan.so (building it first if nvcc is available and it is missing). Raises if it cannot."""
    global _lib
    if _lib is not None:
        return _lib
    path = _build.LIB_PATH
    if not os.path.exists(path):
        try:
            _build.build()
        except Exception as e:  # noqa: BLE001
            raise B2Error("libb2srgan.so is missing and could not be built: %s" % e)
    lib = ctypes.CDLL(path)
    for name, (res, args) in _SIGNATURES.items():
        fn = getattr(lib, name)  # AttributeError if the symbol is not exported
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, what):
    if rc != 0:
        msg = load().b2_last_error()
        raise B2Error("%s failed: %s" % (what, msg.decode(errors="replace") if msg else rc))


def ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


def stream_ptr():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def make_epilogue(out, scale=None, shift=None, alpha=None, act=ACT_NONE, leaky=0.0, residual=None, f16=0, gate=None,
                  gate_slope=0.0, stats_part=None):
    import torch
    ep = Epilogue()
    ep.scale = scale.data_ptr() if scale is not None else None
    ep.shift = shift.data_ptr() if shift is not None else None
    ep.alpha = alpha.data_ptr() if alpha is not None else None
    ep.act = act
    ep.leaky = leaky
    ep.residual = residual.data_ptr() if residual is not None else None
    ep.out = out.data_ptr()
    ep.out_fp32 = 1 if out.dtype == torch.float32 else 0
    ep.f16 = int(f16)
    ep.gate = gate.data_ptr() if gate is not None else None
    ep.gate_slope = gate_slope
    ep.stats_part = stats_part.data_ptr() if stats_part is not None else None
    return ep
