"""ctypes binding of libxvector.so (include/xvector.h).  No CPU fallback: importing works without a
GPU (symbols can be inspected), every compute call raises XVectorError when no B200 is present, and a missing
library raises at load time."""
from __future__ import annotations

import ctypes as C
import os

XV_OK = 0
XV_EINVAL, XV_ECUDA, XV_ENOMEM, XV_ESTATE, XV_ENOTFOUND, XV_EUNSUPPORTED, XV_ECAPACITY = -1, -2, -3, -4, -5, -6, -7
DTYPE_F32, DTYPE_BF16 = 0, 1

LIB_PATH = os.path.join(os.path.dirname(os.path.abspath(__file__)), "lib", "libxvector.so")


class XVectorError(RuntimeError):
    def __init__(self, code, message):
        super().__init__("libxvector error %d: %s" % (code, message))
        self.code = code
        self.message = message


_p = C.c_void_p
_i32p = C.POINTER(C.c_int32)
_i64p = C.POINTER(C.c_int64)
_f32p = C.POINTER(C.c_float)

# name -> (restype, argtypes): every symbol include/xvector.h declares
PROTOTYPES = {
    "xv_create": (C.c_int, [C.c_int, C.c_char_p, C.POINTER(_p)]),
    "xv_destroy": (None, [_p]),
    "xv_config_check": (C.c_int, [C.c_char_p, C.c_char_p, C.c_size_t]),
    "xv_last_error": (C.c_char_p, [_p]),
    "xv_set_weight": (C.c_int, [_p, C.c_char_p, _f32p, _i64p, C.c_int]),
    "xv_finalize_weights": (C.c_int, [_p]),
    "xv_node_info": (C.c_int, [_p, C.c_char_p, _i32p, _i32p, _i32p]),
    "xv_forward": (C.c_int, [_p, _p, _i32p, C.c_int32, C.c_char_p, _f32p, C.c_int64]),
    "xv_forward_frames": (C.c_int, [_p, _p, _i32p, C.c_int32, C.c_char_p, _f32p, _i64p, C.c_int64]),
    "xv_forward_views": (C.c_int, [_p, _p, C.c_int64, _i64p, _i32p, C.c_int32, C.c_char_p, _f32p, C.c_int64]),
    "xv_stage": (C.c_int, [_p, _p, _i32p, C.c_int32]),
    "xv_stage_views": (C.c_int, [_p, _p, C.c_int64, _i64p, _i32p, C.c_int32]),
    "xv_run": (C.c_int, [_p, C.c_char_p]),
    "xv_fetch": (C.c_int, [_p, _f32p, C.c_int64]),
    "xv_sync": (C.c_int, [_p]),
    "xv_stat_pooling": (C.c_int, [_p, _p, _i32p, C.c_int32, C.c_int32, C.c_int, _f32p]),
    "xv_stat_pooling_dev": (C.c_int, [_p, _p, _p, C.c_int32, C.c_int32, C.c_int, C.c_int32, _p]),
    "xv_stat_pooling_noalloc": (C.c_int, [_p, _p, _p, C.c_int32, C.c_int32, C.c_int, C.c_int64, _p]),
    "xv_attention_pooling": (C.c_int, [_p, _p, _p, _f32p, _i32p, C.c_int32, C.c_int32, C.c_int32, C.c_int32,
                                       C.c_int, C.c_int, C.c_int, C.c_int, _f32p, _f32p]),
    "xv_dense_rows_f32": (C.c_int, [_p, _f32p, C.c_int64, C.c_int32, C.c_int32, _f32p, C.c_int32, _f32p, _f32p,
                                    _f32p, _f32p, C.c_int, _f32p]),
    "xv_prelu_f32": (C.c_int, [_p, _f32p, C.c_int64, C.c_int32, _f32p, C.c_int, _f32p]),
    "xv_l2_scaling_f32": (C.c_int, [_p, _f32p, C.c_int64, C.c_int32, C.c_float, C.c_float, _f32p]),
    "xv_event_record": (C.c_int, [_p, C.c_int]),
    "xv_event_elapsed_ms": (C.c_int, [_p, C.c_int, C.c_int, _f32p]),
    "xv_launch_count": (C.c_int64, [_p]),
    "xv_launch_summary": (C.c_int, [_p, C.c_char_p, C.c_size_t]),
    "xv_profile_enable": (C.c_int, [_p, C.c_int]),
    "xv_profile_query": (C.c_int, [_p, C.c_char_p, _f32p, _i64p]),
    "xv_host_alloc": (_p, [C.c_size_t]),
    "xv_host_free": (None, [_p]),
    "xv_device_alloc": (_p, [_p, C.c_size_t]),
    "xv_device_free": (None, [_p, _p]),
    "xv_memcpy_h2d": (C.c_int, [_p, _p, _p, C.c_size_t]),
    "xv_memcpy_d2h": (C.c_int, [_p, _p, _p, C.c_size_t]),
    "xv_upload_bf16": (C.c_int, [_p, _p, _f32p, C.c_size_t]),
    "xv_flush_l2": (C.c_int, [_p]),
    "xv_gemm_bf16": (C.c_int, [_p, _p, C.c_int64, C.c_int32, C.c_int32, C.c_int32, _p, C.c_int32, _p, _p, C.c_float,
                               _p, C.c_int32, C.c_int64]),
    "xv_gemm_bf16_async": (C.c_int, [_p, _p, C.c_int64, C.c_int32, C.c_int32, C.c_int32, _p, C.c_int32, _p, _p, _p,
                                     _p, C.c_int64]),
    "xv_conv2d_bf16_async": (C.c_int, [_p, _p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, _p, C.c_int32, C.c_int32,
                                       C.c_int32, C.c_int32, C.c_int32, _p, _p, _p, _p, _p]),
    "xv_wgrad_bf16_async": (C.c_int, [_p, _p, C.c_int64, C.c_int32, _p, C.c_int32, C.c_int64, C.c_int32, C.c_int32,
                                      C.c_int32, _p, C.c_int32]),
    "xv_train_init": (C.c_int, [_p, C.c_int32]),
    "xv_train_step": (C.c_int, [_p, _p, _p, C.c_int32, C.c_int32, C.c_float, C.c_int64, _p]),
    "xv_train_submit": (C.c_int, [_p, _p, _p, C.c_int32, C.c_int32, C.c_float, C.c_int64]),
    "xv_train_collect": (C.c_int, [_p, _f32p]),
    "xv_train_get_weight": (C.c_int, [_p, C.c_char_p, _p, C.c_int64]),
    "xv_train_get_grad": (C.c_int, [_p, C.c_char_p, _p, C.c_int64]),
    "xv_train_num_variables": (C.c_int, [_p]),
    "xv_train_variable_name": (C.c_int, [_p, C.c_int32, C.c_char_p, C.c_size_t, _i64p]),
    "xv_comm_info": (C.c_int, [_p, _i32p, _i32p, _i64p, _i32p]),
    "xv_comm_bench": (C.c_int, [_p, C.c_int32, _f32p]),
    "xv_comm_dtype": (C.c_int, [_p]),
    "xv_comm_unique_id": (C.c_int, [_p, _p]),
    "xv_comm_init": (C.c_int, [_p, _p, C.c_int, C.c_int]),
    "xv_vec_normalize_length": (C.c_int, [_p, _f32p, C.c_int64, C.c_int32, C.c_int, _f32p]),
    "xv_vec_group_mean": (C.c_int, [_p, _f32p, C.c_int64, C.c_int32, _i32p, C.c_int32, _f32p, _i32p]),
    "xv_cosine_scores": (C.c_int, [_p, _f32p, C.c_int64, _f32p, C.c_int64, C.c_int32, C.c_float, _f32p]),
    "xv_version": (C.c_char_p, []),
}

_lib = None


def load():
    """Loads libxvector.so once.  Raises RuntimeError (loudly) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.isfile(LIB_PATH):
        raise RuntimeError(
            "libxvector.so not found at %s -- build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "or `make -C tf-kaldi-speaker_b200/csrc`. There is no CPU fallback." % LIB_PATH)
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in PROTOTYPES.items():
        fn = getattr(lib, name)          # AttributeError here = header/library mismatch
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib


def check(rc, ctx=None):
    if rc != XV_OK:
        msg = load().xv_last_error(ctx)
        raise XVectorError(rc, msg.decode("utf-8", "replace") if msg else "")
    return rc
