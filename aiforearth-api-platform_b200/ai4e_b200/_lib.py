"""ctypes binding of libai4e_b200.so (the C ABI declared in include/ai4e_b200.h).

There is no CPU fallback: if the shared library is missing or fails to load, importing the
compute API raises `Ai4eLibraryError` with the build command to run.
"""
from __future__ import annotations

import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("AI4E_B200_LIB") or os.path.join(_HERE, "libai4e_b200.so")  # env: A/B builds

OK = 0
ERR_INVALID, ERR_CUDA, ERR_UNSUPPORTED, ERR_NOMEM, ERR_STATE, ERR_TIMEOUT = -1, -2, -3, -4, -5, -6
LAYOUT_NCHW, LAYOUT_NHWC = 0, 1
CONV_AUTO, CONV_IGEMM, CONV_DIRECT = 0, 1, 2
OP_CONV, OP_MAXPOOL, OP_HEAD = 0, 1, 2
TOPK = 5
HEAD_NOALLOC_MAX_FLOATS = 64 * 2048
IMAGE_UNKNOWN, IMAGE_JPEG, IMAGE_PNG = 0, 1, 2


class Ai4eLibraryError(RuntimeError):
    pass


class Ai4eError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"ai4e_b200 error {code}: {msg}")
        self.code = code


class LayerDesc(C.Structure):
    _fields_ = [("op", C.c_int32), ("in_buf", C.c_int32), ("out_buf", C.c_int32), ("res_buf", C.c_int32),
                ("cin", C.c_int32), ("cout", C.c_int32), ("kh", C.c_int32), ("kw", C.c_int32),
                ("stride", C.c_int32), ("pad", C.c_int32), ("relu", C.c_int32), ("reserved", C.c_int32),
                ("w_off", C.c_int64), ("b_off", C.c_int64)]


class ModelDesc(C.Structure):
    _fields_ = [("n_layers", C.c_int32), ("n_bufs", C.c_int32), ("in_h", C.c_int32), ("in_w", C.c_int32),
                ("in_c", C.c_int32), ("in_c_pad", C.c_int32), ("num_classes", C.c_int32),
                ("reserved", C.c_int32), ("mean", C.c_float * 4), ("std", C.c_float * 4),
                ("layers", C.POINTER(LayerDesc))]


class JpegInfo(C.Structure):
    """ai4e_jpeg_info of include/ai4e_b200.h"""
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("components", C.c_int32), ("hmax", C.c_int32),
                ("vmax", C.c_int32), ("h", C.c_int32 * 3), ("v", C.c_int32 * 3), ("blocks_w", C.c_int32 * 3),
                ("blocks_h", C.c_int32 * 3), ("comp_w", C.c_int32 * 3), ("comp_h", C.c_int32 * 3),
                ("restart_interval", C.c_int32), ("coef_off", C.c_int64 * 3), ("record_int16", C.c_int64),
                ("plane_bytes", C.c_int64)]


class EngineStats(C.Structure):
    _fields_ = [("images", C.c_uint64), ("batches", C.c_uint64), ("batch_hist", C.c_uint64 * 11),
                ("queue_depth", C.c_uint64), ("gpu_ms_total", C.c_double), ("lat_ms_p50", C.c_double),
                ("lat_ms_p99", C.c_double), ("healthy", C.c_int32), ("launches_per_forward", C.c_int32)]


_vp, _i, _i64, _u64 = C.c_void_p, C.c_int, C.c_int64, C.c_uint64
_fp = C.POINTER(C.c_float)
_ip = C.POINTER(C.c_int32)
_u64p = C.POINTER(C.c_uint64)

# name -> (restype, argtypes); must list every symbol include/ai4e_b200.h declares
SIGNATURES = {
    "ai4e_version": (_i, []),
    "ai4e_last_error": (C.c_char_p, []),
    "ai4e_device_count": (_i, [C.POINTER(C.c_int)]),
    "ai4e_preprocess": (_i, [_vp, _i, _i, _i, _i, _i64, _i, _i, _i, _i, _i, _fp, _fp, _i, _i, _vp, _vp]),
    "ai4e_preprocess_tile": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _i, _fp, _fp, _i, _i, _vp, _vp]),
    "ai4e_conv_layer": (_i, [_vp, _i, _i, _i, _i, _vp, _vp, _i, _i, _i, _i, _i, _vp, _i, _vp, _i, _vp]),
    "ai4e_jpeg_entropy_decode": (_i, [_vp, _u64, _vp, _u64, C.POINTER(JpegInfo)]),
    "ai4e_jpeg_reconstruct": (_i, [_vp, _i, C.POINTER(JpegInfo), _vp, _vp, _vp]),
    "ai4e_jpeg_decode": (_i, [_vp, _u64, _vp, _u64, C.POINTER(C.c_int32), C.POINTER(C.c_int32), _vp]),
    "ai4e_png_decode": (_i, [_vp, _u64, _vp, _u64, C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "ai4e_image_probe": (_i, [_vp, _u64, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "ai4e_bottleneck_tail": (_i, [_vp, _i, _i, _i, _i, _vp, _vp, _i, _i, _i, _i, _i, _vp, _vp, _i, _vp, _vp, _vp]),
    "ai4e_stem_pool": (_i, [_vp, _i, _i, _i, _vp, _vp, _i, _vp, _vp]),
    "ai4e_preprocess_stem_pool": (_i, [_vp, _i, _i, _i, _fp, _fp, _vp, _vp, _i, _vp, _vp]),
    "ai4e_maxpool": (_i, [_vp, _i, _i, _i, _i, _i, _i, _i, _vp, _vp]),
    "ai4e_head": (_i, [_vp, _i, _i, _i, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp]),
    "ai4e_head_noalloc": (_i, [_vp, _i, _i, _i, _vp, _vp, _i, _vp, _vp, _vp, _i, _vp]),
    "ai4e_engine_create": (_i, [_i, C.POINTER(ModelDesc), _vp, _u64, _i, _i, C.POINTER(_vp)]),
    "ai4e_engine_destroy": (_i, [_vp]),
    "ai4e_engine_forward_host": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp]),
    "ai4e_engine_forward_device": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp]),
    "ai4e_engine_submit": (_i, [_vp, _vp, _i, _i, _i, _i, _u64, _u64p]),
    "ai4e_engine_submit_jpeg": (_i, [_vp, _vp, _u64, _u64, C.POINTER(_u64)]),
    "ai4e_engine_submit_image": (_i, [_vp, _vp, _u64, _u64, C.POINTER(_u64)]),
    "ai4e_pool_submit_image": (_i, [_vp, _vp, _u64, _u64, _u64p]),
    "ai4e_engine_submit_batch": (_i, [_vp, _vp, _i, _i, _i, _i, _u64p]),
    "ai4e_engine_profile": (_i, [_vp, _vp, _i, _i, _i, _i, _fp, _i]),
    "ai4e_engine_wait": (_i, [_vp, _u64, _vp, _vp, _vp]),
    "ai4e_engine_flush": (_i, [_vp]),
    "ai4e_engine_stats_get": (_i, [_vp, C.POINTER(EngineStats)]),
    "ai4e_engine_read_activation": (_i, [_vp, _i, _vp, _u64]),
    "ai4e_pool_create": (_i, [C.POINTER(C.c_int), _i, C.POINTER(ModelDesc), _vp, _u64, _i, _i, C.POINTER(_vp)]),
    "ai4e_pool_destroy": (_i, [_vp]),
    "ai4e_pool_submit": (_i, [_vp, _vp, _i, _i, _i, _i, _u64, _u64p]),
    "ai4e_pool_submit_jpeg": (_i, [_vp, _vp, _u64, _u64, _u64p]),
    "ai4e_engine_submit_tile": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp, _i, C.POINTER(C.c_int32)]),
    "ai4e_engine_submit_tile_summary": (_i, [_vp, _vp, _i, _i, _i, _i, _i, C.POINTER(C.c_uint64), C.POINTER(C.c_int32)]),
    "ai4e_engine_wait_tile_summary": (_i, [_vp, C.c_uint64, _vp, _vp, _i, _vp, _vp, C.POINTER(C.c_int32)]),
    "ai4e_tile_summary": (_i, [_vp, _vp, _i, _i, _i, _vp, _vp, _vp, _vp, _vp]),
    "ai4e_engine_wait_many": (_i, [_vp, _vp, _i, _vp, _vp]),
    "ai4e_pool_wait": (_i, [_vp, _u64, _vp, _vp, _vp]),
    "ai4e_engine_submit_many": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp, _i, _vp]),
    "ai4e_engine_poll": (_i, [_vp, _i, _i64, _vp, _vp, _vp, _vp, _vp, C.POINTER(C.c_int)]),
    "ai4e_engine_wait_for": (_i, [_vp, _u64, _i64, _vp, _vp, _vp]),
    "ai4e_engine_debug_inject_fault": (_i, [_vp, _i]),
    "ai4e_pool_submit_many": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp, _i, _vp]),
    "ai4e_pool_submit_batch": (_i, [_vp, _vp, _i, _i, _i, _i, _u64p]),
    "ai4e_pool_openloop": (_i, [_vp, _vp, _i, _i, _i, _i, C.c_double, C.c_double, _i, C.c_uint64, _vp, C.c_int64,
                                C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_int64), C.POINTER(C.c_double),
                                C.POINTER(C.c_double)]),
    "ai4e_pool_submit_tile": (_i, [_vp, _vp, _i, _i, _i, _i, _i, _vp, _i, C.POINTER(C.c_int32)]),
    "ai4e_pool_submit_tile_summary": (_i, [_vp, _vp, _i, _i, _i, _i, _i, C.POINTER(C.c_uint64), C.POINTER(C.c_int32)]),
    "ai4e_pool_wait_tile_summary": (_i, [_vp, C.c_uint64, _vp, _vp, _i, _vp, _vp, C.POINTER(C.c_int32)]),
    "ai4e_pool_wait_many": (_i, [_vp, _vp, _i, _vp, _vp]),
    "ai4e_pool_poll": (_i, [_vp, _i, _i64, _vp, _vp, _vp, _vp, _vp, C.POINTER(C.c_int)]),
    "ai4e_pool_wait_for": (_i, [_vp, _u64, _i64, _vp, _vp, _vp]),
    "ai4e_pool_flush": (_i, [_vp]),
    "ai4e_pool_size": (_i, [_vp]),
    "ai4e_pool_engine": (_i, [_vp, _i, C.POINTER(_vp)]),
}

_lib = None


def load() -> C.CDLL:
    """Load the shared library (once) and type every entry point.  Fails loudly."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise Ai4eLibraryError(
            f"{LIB_PATH} is missing: build it with `python aiforearth-api-platform_b200/build.py` "
            "(nvcc, sm_100a).  There is no CPU fallback.")
    try:
        lib = C.CDLL(LIB_PATH)
    except OSError as e:  # pragma: no cover - depends on the machine
        raise Ai4eLibraryError(f"cannot load {LIB_PATH}: {e}") from e
    for name, (res, args) in SIGNATURES.items():
        try:
            fn = getattr(lib, name)
        except AttributeError as e:
            if os.environ.get("AI4E_B200_DEV_PARTIAL") == "1":  # bring-up of a half-built library only
                continue
            raise Ai4eLibraryError(f"{LIB_PATH} does not export {name}: rebuild it") from e
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc: int) -> None:
    if rc != OK:
        msg = load().ai4e_last_error()
        raise Ai4eError(rc, msg.decode("utf-8", "replace") if msg else "")
