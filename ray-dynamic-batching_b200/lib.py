"""ctypes binding of libb200infer.so (the C ABI declared in include/b200infer.h).

The library is the product: there is no Python/torch fallback.  If it cannot be built or loaded, `load()` raises.
"""
import ctypes
import os
import threading

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB_PATH = os.path.join(_HERE, "libb200infer.so")
_lock = threading.Lock()
_lib = None

c_void_p, c_int, c_float, c_size_t, c_char_p = (ctypes.c_void_p, ctypes.c_int, ctypes.c_float, ctypes.c_size_t,
                                                ctypes.c_char_p)
c_float_p = ctypes.POINTER(ctypes.c_float)
c_int_p = ctypes.POINTER(ctypes.c_int)

# name -> (restype, argtypes); every symbol include/b200infer.h declares must be listed (tests check both ways)
SIGNATURES = {
    "b200_last_error": (c_char_p, []),
    "b200_abi_version": (c_int, []),
    "b200_conv2d_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int,
                                 c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "b200_conv1x1_dual_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_int, c_int, c_int, c_int,
                                       c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_void_p]),
    "b200_conv1x1_shuffle_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_int, c_int, c_int,
                                          c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "b200_conv1x1_scaled_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_void_p,
                                         c_int, c_void_p, c_int, c_int, c_int, c_void_p]),
    "b200_gather_conv_npad": (c_int, [c_int]),
    "b200_gather_conv_kpad": (c_int, [c_int, c_int]),
    "b200_gather_conv2d_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int,
                                        c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p]),
    "b200_stem_cout_tile": (c_int, [c_int]),
    "b200_stem_conv_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                    c_int, c_int, c_int, c_void_p]),
    "b200_stem_row_weight_bytes": (c_int, []),
    "b200_stem_row_conv_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "b200_stem_pool_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int, c_void_p]),
    "b200_stem_row_weight_bytes_k": (c_int, [c_int, c_int]),
    "b200_stem_row_conv_k_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                          c_int, c_void_p]),
    "b200_stem_pool_k_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                      c_int, c_void_p]),
    "b200_stem_tc_kpad": (c_int, [c_int]),
    "b200_stem_tc_nout": (c_int, [c_int]),
    "b200_stem_conv_tc_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int,
                                       c_int, c_int, c_int, c_void_p]),
    "b200_dwconv3x3_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int,
                                    c_int, c_int, c_void_p]),
    "b200_dwconv3x3_strips": (c_int, [c_int, c_int, c_int]),
    "b200_dwconv3x3_squeeze_rows": (c_int, [c_int, c_int, c_int, c_int]),
    "b200_dwconv3x3_squeeze_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p,
                                            c_int, c_int, c_int, c_void_p, c_void_p]),
    "b200_se_gate": (c_int, [c_void_p, c_int, c_int, c_int, c_float, c_void_p, c_void_p, c_void_p, c_void_p, c_int,
                             c_void_p, c_void_p]),
    "b200_se_hidden": (c_int, [c_void_p, c_int, c_int, c_int, c_float, c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "b200_se_scale_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_void_p, c_int,
                                   c_void_p, c_int, c_void_p]),
    "b200_maxpool_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_int, c_int, c_int, c_int,
                                  c_void_p]),
    "b200_global_avgpool_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p]),
    "b200_fc": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_float, c_int, c_void_p]),
    "b200_channel_scale_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_void_p, c_int,
                                        c_void_p]),
    "b200_copy_channels_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_int, c_void_p]),
    "b200_upsample2x_nhwc": (c_int, [c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_int, c_void_p]),
    "b200_shuffle_concat_nhwc": (c_int, [c_void_p, c_int, c_void_p, c_int, c_int, c_int, c_int, c_int, c_void_p, c_int,
                                         c_void_p]),
    "b200_engine_create": (c_int, [c_char_p, c_void_p, c_size_t, c_int, c_int, ctypes.POINTER(c_void_p)]),
    "b200_engine_create_ex": (c_int, [c_char_p, c_void_p, c_size_t, c_int, c_int, ctypes.c_uint, ctypes.POINTER(c_void_p)]),
    "b200_engine_warmup": (c_int, [c_void_p, c_void_p, c_int, c_void_p]),
    "b200_engine_memory": (c_int, [c_void_p, ctypes.POINTER(c_size_t), ctypes.POINTER(c_size_t)]),
    "b200_plan_memory": (c_int, [c_void_p, c_size_t, c_int, ctypes.c_uint, ctypes.POINTER(c_size_t)]),
    "b200_engine_destroy": (None, [c_void_p]),
    "b200_engine_info": (c_int, [c_void_p, c_int_p, c_int_p, c_int_p, c_int_p, c_int_p]),
    "b200_engine_set_graphs": (c_int, [c_void_p, c_int]),
    "b200_engine_launch_count": (ctypes.c_longlong, [c_void_p]),
    "b200_engine_buffer": (c_int, [c_void_p, c_int, ctypes.POINTER(c_void_p), ctypes.POINTER(c_size_t)]),
    "b200_engine_profile": (c_int, [c_void_p, c_int, c_void_p, c_void_p, c_int, c_void_p]),
    "b200_engine_read_buffer": (c_int, [c_void_p, c_int, c_int, c_void_p]),
    "b200_collate": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_size_t, c_void_p,
                             c_void_p]),
    "b200_collate_u8_noalloc": (c_int, [c_void_p, c_void_p, c_void_p, c_int, c_int, c_int, c_int, c_int, c_int, c_void_p,
                                        c_void_p, c_size_t, c_void_p, c_void_p]),
    "b200_collate_f32": (c_int, [c_void_p, c_void_p, c_int, c_int, c_int, c_void_p, c_void_p]),
    "b200_decode_yolo": (c_int, [c_void_p, c_void_p, c_int, c_float, c_float, c_int, c_void_p, c_void_p, c_void_p]),
    "b200_infer": (c_int, [c_void_p, c_void_p, c_int, c_void_p, c_void_p]),
    "b200_ring_bytes": (c_size_t, [ctypes.c_uint32, ctypes.c_uint32]),
    "b200_ring_init": (c_int, [c_void_p, c_size_t, ctypes.c_uint32, ctypes.c_uint32, ctypes.c_uint32]),
    "b200_ring_layout": (c_int, [c_void_p, c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "b200_ring_push": (c_int, [c_void_p, c_char_p, c_void_p, ctypes.c_uint32, c_int, c_int, c_int, ctypes.c_double]),
    "b200_ring_push_many": (c_int, [c_void_p, c_int, c_void_p, c_char_p, ctypes.c_longlong, c_void_p, c_size_t,
                                    ctypes.c_uint32, c_int, c_int, c_int, ctypes.c_double]),
    "b200_ring_pop": (c_int, [c_void_p, c_int, ctypes.c_double, ctypes.c_double, ctypes.c_double, c_void_p, c_void_p,
                              c_void_p, c_void_p, c_void_p, c_void_p, c_void_p]),
    "b200_ring_release": (c_int, [c_void_p, c_void_p, c_int]),
    "b200_ring_qsize": (ctypes.c_longlong, [c_void_p]),
    "b200_ring_record": (c_int, [c_void_p, ctypes.c_double, c_int, ctypes.c_double]),
    "b200_ring_stats": (c_int, [c_void_p, c_void_p, c_int]),
    "b200_ring_latencies": (c_int, [c_void_p, c_void_p, c_int]),
}


ENGINE_KEEP_BUFFERS = 1   # B200_ENGINE_KEEP_BUFFERS in include/b200infer.h


class B200Error(RuntimeError):
    pass


def lib_path():
    return _LIB_PATH


def load(build_if_missing=True):
    """Returns the loaded CDLL; builds it first when the sources are newer or the .so is missing (needs nvcc)."""
    global _lib
    with _lock:
        if _lib is not None:
            return _lib
        if build_if_missing:
            from . import build as _build
            _build.build()
        if not os.path.exists(_LIB_PATH):
            raise B200Error("libb200infer.so is missing (%s): run `python ray-dynamic-batching_b200/build.py`; "
                            "there is no fallback path" % _LIB_PATH)
        lib = ctypes.CDLL(_LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            fn = getattr(lib, name)  # AttributeError if the symbol is missing: fail loudly
            fn.restype = res
            fn.argtypes = args
        if lib.b200_abi_version() != 1:
            raise B200Error("libb200infer ABI version mismatch")
        _lib = lib
        return _lib


def check(rc, what=""):
    if rc != 0:
        msg = load().b200_last_error()
        raise B200Error("%s failed (rc=%d): %s" % (what or "libb200infer call", rc, msg.decode() if msg else "?"))


def ptr(t):
    """Device/host pointer of a torch tensor (or None) as c_void_p."""
    return None if t is None else ctypes.c_void_p(t.data_ptr())


def stream_ptr():
    import torch
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
