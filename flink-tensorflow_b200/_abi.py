"""ctypes binding of include/ftf.h (the C ABI of libftf_b200.so).

The library is the product: if it is missing this module raises, there is no Python or CPU path behind it.
"""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("FTF_LIB_PATH") or os.path.join(_HERE, "libftf_b200.so")   # the override serves A/B measurements of two builds

OK, INVALID_ARGUMENT, NOT_FOUND, RESOURCE_EXHAUSTED = 0, 3, 5, 8
FAILED_PRECONDITION, OUT_OF_RANGE, UNIMPLEMENTED, INTERNAL, VERSION_MISMATCH = 9, 11, 12, 13, 100

FLOAT, DOUBLE, INT32, UINT8, STRING, INT64, BOOL, BFLOAT16 = 1, 2, 3, 4, 7, 9, 10, 14


class FtfTensor(C.Structure):
    _fields_ = [("dtype", C.c_int32), ("rank", C.c_int32), ("dims", C.POINTER(C.c_int64)),
                ("data", C.c_void_p), ("nbytes", C.c_int64)]


class FtfConfig(C.Structure):
    _fields_ = [("device", C.c_int32), ("max_batch", C.c_int32), ("max_delay_us", C.c_int32),
                ("out_h", C.c_int32), ("out_w", C.c_int32), ("mean", C.c_float), ("scale", C.c_float),
                ("compute_dtype", C.c_int32), ("top_k", C.c_int32), ("num_classes", C.c_int32),
                ("seed", C.c_uint64)]


class FtfStats(C.Structure):
    _fields_ = [("records", C.c_uint64), ("batches", C.c_uint64), ("runs", C.c_uint64),
                ("h2d_bytes", C.c_uint64), ("d2h_bytes", C.c_uint64), ("kernel_launches", C.c_uint64),
                ("last_run_device_ms", C.c_double), ("queue_latency_us_p50", C.c_double),
                ("queue_latency_us_p99", C.c_double)]


# name -> (restype, argtypes); every symbol include/ftf.h declares
SYMBOLS = {
    "ftf_default_config": (None, [C.POINTER(FtfConfig)]),
    "ftf_create": (C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.POINTER(FtfConfig), C.POINTER(C.c_void_p)]),
    "ftf_destroy": (C.c_int, [C.c_void_p]),
    "ftf_run": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(C.c_char_p), C.POINTER(FtfTensor), C.c_int32,
                          C.POINTER(C.c_char_p), C.POINTER(FtfTensor)]),
    "ftf_release": (None, [C.c_void_p, C.POINTER(FtfTensor), C.c_int32]),
    "ftf_submit": (C.c_int, [C.c_void_p, C.c_uint64, C.POINTER(FtfTensor)]),
    "ftf_poll": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(C.c_uint64), C.POINTER(FtfTensor), C.c_int32,
                           C.POINTER(C.c_int32)]),
    "ftf_flush": (C.c_int, [C.c_void_p]),
    "ftf_poll_status": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.c_int32, C.POINTER(C.c_char_p)]),
    "ftf_poll_select": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(C.c_char_p)]),
    "ftf_last_error": (C.c_char_p, [C.c_void_p]),
    "ftf_stats_get": (C.c_int, [C.c_void_p, C.POINTER(FtfStats)]),
    "ftf_signature_info": (C.c_char_p, [C.c_void_p, C.c_int32, C.c_int32]),
    "ftf_savedmodel_describe": (C.c_int, [C.c_char_p, C.c_char_p, C.c_char_p, C.c_int64, C.POINTER(C.c_int64)]),
    "ftf_host_alloc": (C.c_void_p, [C.c_int64]),
    "ftf_host_free": (None, [C.c_void_p]),
    "ftf_tv_binary_length": (C.c_int64, [C.c_int32, C.c_int64]),
    "ftf_tv_encode": (C.c_int, [C.POINTER(FtfTensor), C.c_void_p, C.c_int64, C.POINTER(C.c_int64)]),
    "ftf_tv_decode": (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(FtfTensor), C.POINTER(C.c_int64)]),
    "ftf_tensor_name_parse": (C.c_int, [C.c_char_p, C.c_char_p, C.c_int32, C.POINTER(C.c_int32)]),
    "ftf_string_tensor_encode": (C.c_int64, [C.c_int64, C.POINTER(C.c_char_p), C.POINTER(C.c_int64), C.c_void_p,
                                             C.c_int64]),
    "ftf_parse_examples_f32": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, C.c_char_p, C.c_int32, C.c_float,
                                         C.c_void_p]),
    "ftf_affine_f32_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_float, C.c_float, C.c_void_p]),
    "ftf_resize_norm_dev": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_float,
                                      C.c_float, C.c_void_p, C.c_int32, C.c_void_p]),
    "ftf_softmax_topk_dev": (C.c_int, [C.c_void_p, C.c_int64, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p,
                                       C.c_void_p, C.c_void_p, C.c_void_p]),
    "ftf_conv2d_bf16_dev": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p,
                                      C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p,
                                      C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "ftf_maxpool3x3s2_bf16_dev": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int32,
                                            C.c_int32, C.c_void_p]),
    "ftf_pool_branch_bf16_dev": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p,
                                           C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "ftf_pool_branch_bf16_static_dev": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_int32, C.c_void_p,
                                                  C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "ftf_global_avgpool_bf16_dev": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p]),
    "ftf_label_export_weights": (C.c_int, [C.c_void_p, C.c_char_p]),
    "ftf_debug_conv_profile": (C.c_int, [C.POINTER(C.c_uint64), C.c_int32]),
    "ftf_debug_activation": (C.c_int, [C.c_void_p, C.c_char_p, C.c_int64, C.c_void_p, C.c_int64, C.POINTER(C.c_int32)]),
    "ftf_label_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int32, C.c_int64, C.c_void_p, C.c_void_p, C.c_void_p,
                                C.c_void_p]),
    "ftf_jpeg_info": (C.c_int, [C.c_void_p, C.c_int64, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_int32)]),
    "ftf_jpeg_batch_create": (C.c_int, [C.c_int32, C.POINTER(C.c_void_p)]),
    "ftf_jpeg_batch_destroy": (None, [C.c_void_p]),
    "ftf_jpeg_batch_error": (C.c_char_p, [C.c_void_p]),
    "ftf_jpeg_batch_load": (C.c_int, [C.c_void_p, C.c_int32, C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.POINTER(C.c_int64),
                                      C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.c_void_p]),
    "ftf_jpeg_batch_rgb_bytes": (C.c_int64, [C.c_void_p]),
    "ftf_jpeg_batch_scan_bytes": (C.c_int64, [C.c_void_p]),
    "ftf_jpeg_batch_segments": (C.c_int32, [C.c_void_p]),
    "ftf_jpeg_batch_decode_dev": (C.c_int, [C.c_void_p, C.c_void_p, C.c_int64, C.c_void_p]),
    "ftf_jpeg_batch_status": (C.c_int, [C.c_void_p, C.POINTER(C.c_int32), C.c_int32]),
    "ftf_runner_stream": (C.c_void_p, [C.c_void_p]),
    "ftf_launch_count": (C.c_uint64, []),
    "ftf_version": (C.c_char_p, []),
}

_lib = None


def load_library():
    """Loads libftf_b200.so; raises (never falls back) when it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise ImportError(
            f"{LIB_PATH} is missing: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(make -C flink-tensorflow_b200/csrc).  There is no fallback path.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in SYMBOLS.items():
        fn = getattr(lib, name)  # AttributeError if the ABI lost a symbol
        fn.restype = res
        fn.argtypes = args
    _lib = lib
    return lib
