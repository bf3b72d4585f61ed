"""ctypes binding of libb2v.so (include/b2v.h). There is no CPU fallback: if the library is missing, or the device is
not sm_100, calls raise."""
import ctypes as C
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
# B2V_LIB: A/B a differently built libb2v (tools/gemm_ab.py); the default is the in-tree build
LIB_PATH = os.environ.get("B2V_LIB") or os.path.join(_HERE, "libb2v.so")

F32, F16, BF16 = 0, 1, 2
EPI_GEGLU, EPI_GELU = 1, 2

_lock = threading.Lock()
_lib = None


class B2VError(RuntimeError):
    pass


class UNet3DConfigC(C.Structure):
    _fields_ = [
        ("in_channels", C.c_int), ("out_channels", C.c_int), ("block_out_channels", C.c_int * 4),
        ("layers_per_block", C.c_int), ("cross_attention_dim", C.c_int), ("attention_head_dim", C.c_int),
        ("norm_num_groups", C.c_int), ("norm_eps", C.c_float), ("transformer_in_heads", C.c_int),
        ("text_len", C.c_int), ("act_dtype", C.c_int),
    ]


_vp, _i, _i64, _f, _sz, _u64 = C.c_void_p, C.c_int, C.c_int64, C.c_float, C.c_size_t, C.c_uint64

_SIGNATURES = {
    "b2v_version": (_i, []),
    "b2v_last_error": (C.c_char_p, []),
    "b2v_device_check": (_i, []),
    "b2v_launch_count": (_u64, []),
    "b2v_sha256_leaves": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp, _i, _vp]),
    "b2v_merkle_depth": (_i, [_i]),
    "b2v_merkle_scratch_bytes": (_sz, [_i]),
    "b2v_merkle_build": (_i, [_vp, _i, _vp, _vp, _vp, _vp]),
    "b2v_merkle_static_max_leaves": (_i, []),
    "b2v_merkle_build_static": (_i, [_vp, _i, _vp, _vp, _vp]),
    "b2v_merkle_verify": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp]),
    "b2v_merkle_verify_roots": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp]),
    "b2v_base64_len": (_u64, [_u64]),
    "b2v_base64_rows": (_i, [_vp, _i64, _u64, _u64, _vp, _vp]),
    "b2v_base64_decode_rows": (_i, [_vp, _i64, _u64, _u64, _vp, _u64, _vp, _vp]),
    "b2v_spotcheck_ddim": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _i, _i, _i64, _vp, _vp]),
    "b2v_cfg_ddim_step": (_i, [_vp, _vp, _vp, _f, _i, _f, _f, _f, _i, _vp, _vp, _i, _vp, _vp, _i, _i, _i, _i64, _vp]),
    "b2v_make_unet_input": (_i, [_vp, _vp, _f, _vp, _i, _i, _i, _i, _i64, _vp]),
    "b2v_context_mean": (_i, [_vp, _vp, _i, _i, _i, _i64, _vp]),
    "b2v_blend_overlap": (_i, [_vp, _vp, _vp, _i, _vp, _i, _i, _i64, _i, _vp, _vp, _vp]),
    "b2v_unet3d_create": (_i, [C.POINTER(UNet3DConfigC), C.POINTER(_vp)]),
    "b2v_unet3d_destroy": (None, [_vp]),
    "b2v_unet3d_load_tensor": (_i, [_vp, C.c_char_p, _vp, _i, _i, C.POINTER(_i64)]),
    "b2v_unet3d_finalize": (_i, [_vp]),
    "b2v_unet3d_param_count": (_i64, [_vp]),
    "b2v_unet3d_workspace_bytes": (_sz, [_vp, _i, _i, _i, _i]),
    "b2v_unet3d_forward": (_i, [_vp, _vp, _f, _vp, _vp, _i, _i, _i, _i, _i, _vp, _sz, _vp]),
    "b2v_unet3d_flops": (C.c_double, [_vp, _i, _i, _i, _i]),
    "b2v_op_gemm": (_i, [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _i64, _i, _i, _i, _i64, _i64, _i, _i, _vp]),
    "b2v_op_conv3x3_scratch_bytes": (_sz, [_i, _i, _i, _i, _i, _i]),
    "b2v_op_conv3x3": (_i, [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _i, _i, _i, _i, _i, _i, _i, _vp, _i, _vp]),
    "b2v_op_group_norm": (_i, [_vp, _i, _vp, _i, _vp, _vp, _vp, _i64, _i64, _i, _f, _i, _vp, _i, _vp]),
    "b2v_op_group_norm_scratch_bytes": (_sz, [_i64, _i64, _i]),
    "b2v_row_stats_slots": (_i, [_i64, _i, _i64]),
    "b2v_op_gemm_rowstats": (_i, [_vp, _vp, _vp, _vp, _vp, _i64, _i, _i, _i, _vp, _vp]),
    "b2v_op_gemm_lnin": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _f, _vp, _i64, _i, _i, _i, _i, _vp]),
    "b2v_gn_stats_ok": (_i, [_i64, _i, _i64, _i]),
    "b2v_op_gemm_gnstats": (_i, [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _i64, _i, _i, _i, _i64, _i64, _i, _vp, _i, _vp]),
    "b2v_op_conv3x3_gnstats": (_i, [_vp, _vp, _vp, _vp, _i64, _vp, _vp, _i, _i, _i, _i, _i, _i, _vp, _i, _vp]),
    "b2v_op_group_norm_pre": (_i, [_vp, _i, _vp, _vp, _vp, _i64, _i64, _i, _f, _i, _vp, _vp, _i, _vp]),
    "b2v_op_layer_norm": (_i, [_vp, _vp, _vp, _vp, _i64, _i, _f, _i, _vp]),
    "b2v_op_attention": (_i, [_vp, _i64, _vp, _vp, _i64, _vp, _i64, _i, _i, _i, _i, _i, _f, _i, _vp]),
    "b2v_op_temporal_attention": (_i, [_vp, _i64, _i, _i, _i, _vp, _i64, _i, _i, _i64, _i, _f, _i, _vp]),
    "b2v_op_softmax_rows": (_i, [_vp, _i64, _i, _i64, _f, _vp]),
    "b2v_op_frames_u8": (_i, [_vp, _i64, _i64, _vp, _vp]),
    "b2v_op_upsample2x": (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "b2v_op_latent_rows": (_i, [_vp, _i, _vp, _i, _i, _i, _i64, _i, _i, _vp]),
}

EXPORTED_SYMBOLS = tuple(_SIGNATURES)


def load(build_if_missing=True):
    """Return the ctypes handle of libb2v.so, building it in-tree with nvcc when absent."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if not os.path.exists(LIB_PATH):
            if not build_if_missing:
                raise B2VError("libb2v.so is not built (run __graft_entry__.build()); there is no CPU fallback")
            from . import build as _build
            _build.build()
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in _SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype = res
            fn.argtypes = args
        _lib = lib
        return lib


def check(status):
    if status != 0:
        msg = load().b2v_last_error()
        raise B2VError("libb2v error %d: %s" % (status, msg.decode("utf-8", "replace") if msg else "?"))


def last_error():
    msg = load().b2v_last_error()
    return msg.decode("utf-8", "replace") if msg else ""
