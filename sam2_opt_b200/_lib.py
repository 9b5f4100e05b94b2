"""ctypes binding of the C-ABI library (include/sam2b.h).  Loading fails loudly when the library has not been
built; there is no Python / CPU fallback for any compute entry point."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libsam2b.so")

_vp, _i, _i64, _f = C.c_void_p, C.c_int, C.c_int64, C.c_float

# name -> (restype, argtypes); must list every symbol include/sam2b.h declares (tests/test_abi.py checks that)
SIGNATURES = {
    "sam2b_last_error": (C.c_char_p, []),
    "sam2b_version": (C.c_char_p, []),
    "sam2b_launch_count": (C.c_ulonglong, []),
    "sam2b_create": (_i, [_i, C.POINTER(_vp)]),
    "sam2b_destroy": (_i, [_vp]),
    "sam2b_set_tensor": (_i, [_vp, C.c_char_p, _vp, C.POINTER(_i64), _i]),
    "sam2b_finalize": (_i, [_vp]),
    "sam2b_image_encode": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "sam2b_image_encode_host": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "sam2b_image_encode_host_async": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "sam2b_host_sync": (_i, [_vp, _vp]),
    "sam2b_image_encode_host_u8": (_i, [_vp, _vp, _i, _i, _vp, _vp, _vp, _vp]),
    "sam2b_position_encoding": (_i, [_i, _i, _i, _vp, _vp]),
    "sam2b_memory_attention": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _vp, _vp]),
    "sam2b_memory_attention_varlen": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, C.POINTER(C.c_int32), C.POINTER(C.c_int32), _vp, _vp]),
    "sam2b_memory_attention_varlen_dev": (_i, [_vp, _vp, _vp, _vp, _vp, _i, _vp, _vp, _i, _i, _vp, _vp, _vp]),
    "sam2b_memory_encode": (_i, [_vp, _vp, _vp, _i, _vp, _vp]),
    "sam2b_memory_encode_low_res": (_i, [_vp, _vp, _i64, _vp, _i, _vp, _i, _vp, _vp]),
    "sam2b_sam_heads": (_i, [_vp, _vp, _i64, _i64, _i64, _vp, _i64, _vp, _i64, _vp, _vp, _i, _vp, _i, _vp, _i, _i, _i,
                             _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp, _vp]),
    "sam2b_mask_downscale": (_i, [_vp, _vp, _i, _vp, _vp]),
    "sam2b_stream_acquire": (_i, [_vp, _vp]),
    "sam2b_stream_release": (_i, [_vp, _vp]),
    "sam2b_workspace_generation": (C.c_ulonglong, [_vp]),
    "sam2b_peer_alloc": (_i, [_i, _i64, C.POINTER(_vp), C.c_char_p]),
    "sam2b_peer_free": (_i, [_i, _vp]),
    "sam2b_peer_open": (_i, [_i, C.c_char_p, C.POINTER(_vp)]),
    "sam2b_peer_close": (_i, [_i, _vp]),
    "sam2b_peer_copy": (_i, [_vp, _vp, _i64, _vp]),
    "sam2b_profile_begin": (_i, []),
    "sam2b_profile_end": (_i, [C.c_char_p, _i64, C.POINTER(_i64)]),
    "sam2b_op_gemm": (_i, [_vp, _i64, _i64, _vp, _i64, _i64, _i, _i, _i, _i, _f, _vp, _i, _i, _vp, _vp, _i64, _i64, _i,
                           _vp, _vp, _i64, _i64, _i, _vp]),
    "sam2b_op_gemm_ln_producer": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp, _i, C.POINTER(_i), _i, _vp]),
    "sam2b_op_gemm_ln_consumer": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _i, _f, _i, _vp, _vp, _i, _vp]),
    "sam2b_op_layernorm": (_i, [_vp, _vp, _vp, _f, _vp, _vp, _i64, _i, _vp]),
    "sam2b_op_cast_bf16": (_i, [_vp, _vp, _i64, _vp]),
    "sam2b_op_maxpool2x2_nhwc": (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "sam2b_op_window_attention": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _i, _vp]),
    "sam2b_op_connected_components": (_i, [_vp, _vp, _vp, _i, _i, _i, _vp]),
    "sam2b_op_preprocess_u8": (_i, [_vp, _vp, _i, _i, _i, _i, _vp]),
    "sam2b_op_resize_frames_workspace": (_i64, [_i, _i, _i, _i, _i]),
    "sam2b_op_resize_frames": (_i, [_vp, _i, _i, _i, _i, _vp, _i, _i, _i, _vp, _vp]),
    "sam2b_op_resize_frames_static": (_i, [_vp, _i, _i, _i, _i, _vp, _i, _i, _i, _vp]),
    "sam2b_op_upsample_masks": (_i, [_vp, _vp, _vp, _i64, _i, _i, _i, _i, _f, _vp]),
    "sam2b_op_softmax_rows": (_i, [_vp, _i64, _vp, _i64, _i64, _i, _vp]),
    "sam2b_op_attention_hd256_workspace": (_i64, [_i, _i, _i64]),
    "sam2b_op_attention_hd256_v64_workspace": (_i64, [_i, _i, _i64]),
    "sam2b_op_attention_hd256_v64": (_i, [_vp, _vp, _vp, _i, _i, _i64, _vp, _vp, _vp, _vp]),
    "sam2b_op_attention_hd256": (_i, [_vp, _vp, _vp, _i, _i, _i64, _vp, _vp, _vp]),
}

_lib = None


def load():
    """Return the loaded library (cached)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(
                f"{LIB_PATH} is missing: build it with `python -m sam2_opt_b200.build` "
                "(sam2_opt_b200 has no fallback path without its CUDA library)")
        lib = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)
            fn.restype, fn.argtypes = res, args
        _lib = lib
    return _lib


def check(rc, what=""):
    if rc != 0:
        raise RuntimeError(f"sam2b {what}: {load().sam2b_last_error().decode()}")
