"""CPU oracle (test infrastructure only; see oracle/oracle.c).  ctypes binding of oracle/liboracle.so."""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_LIB = os.path.join(_HERE, "liboracle.so")
_lib = None


def build():
    subprocess.run(["make", "-C", _HERE], check=True, stdout=subprocess.DEVNULL, stderr=subprocess.PIPE)


def lib():
    global _lib
    if _lib is None:
        srcs = [os.path.join(_HERE, f) for f in os.listdir(_HERE) if f.endswith(".c")]
        if not os.path.exists(_LIB) or any(os.path.getmtime(s) > os.path.getmtime(_LIB) for s in srcs):
            build()
        _lib = C.CDLL(_LIB)
        _lib.orc_tv_binary_length.restype = C.c_int64
        _lib.orc_string_tensor_encode.restype = C.c_int64
        _lib.orc_bf16_round.restype = C.c_float
        _lib.orc_bf16_round.argtypes = [C.c_float]
    return _lib


def num_threads():
    return lib().orc_num_threads()


def _p(a):
    return a.ctypes.data_as(C.c_void_p)


def affine_f32(x, a, b, threads=1):
    x = np.ascontiguousarray(x, np.float32)
    y = np.empty_like(x)
    lib().orc_affine_f32(_p(x), C.c_int64(x.size), C.c_float(a), C.c_float(b), _p(y), C.c_int(threads))
    return y


def resize_bilinear_norm_u8(images, out_h, out_w, mean, scale, threads=1):
    images = np.ascontiguousarray(images, np.uint8)
    n, h, w, ch = images.shape
    out = np.empty((n, out_h, out_w, ch), np.float32)
    rc = lib().orc_resize_bilinear_norm_u8(_p(images), C.c_int64(n), h, w, ch, out_h, out_w, C.c_float(mean),
                                           C.c_float(scale), _p(out), C.c_int(threads))
    if rc:
        raise ValueError(f"oracle resize: status {rc}")
    return out


def bf16_round(a):
    """float32 array -> float32 array holding the bf16 (RNE) rounded values."""
    a = np.ascontiguousarray(a, np.float32)
    u = a.view(np.uint32).astype(np.uint64)
    nan = (u & 0x7fffffff) > 0x7f800000
    r = ((u + 0x7fff + ((u >> 16) & 1)) >> 16) << 16
    r = np.where(nan, (u | 0x400000) & 0xffff0000, r).astype(np.uint32)
    return r.view(np.float32).reshape(a.shape)


def bf16_bits(a):
    return (bf16_round(a).view(np.uint32) >> 16).astype(np.uint16)


def top_labels(scores, n_labels, k):
    scores = np.ascontiguousarray(scores, np.float32)
    m, n = scores.shape
    idx = np.empty((m, k), np.int32)
    sc = np.empty((m, k), np.float32)
    rc = lib().orc_top_labels(_p(scores), C.c_int64(m), C.c_int64(n), C.c_int64(n_labels), k, _p(idx), _p(sc))
    if rc:
        raise ValueError(f"oracle top_labels: status {rc}")
    return idx, sc


def softmax_rows(x):
    x = np.ascontiguousarray(x, np.float32)
    p = np.empty_like(x)
    lib().orc_softmax_rows(_p(x), C.c_int64(x.shape[0]), C.c_int64(x.shape[1]), _p(p))
    return p


def tv_encode(dtype, shape, payload):
    dims = (C.c_int64 * max(1, len(shape)))(*shape)
    cap = lib().orc_tv_binary_length(len(shape), C.c_int64(len(payload)))
    out = C.create_string_buffer(max(int(cap), 1))
    rc = lib().orc_tv_encode(dtype, len(shape), dims, payload, C.c_int64(len(payload)), out, C.c_int64(cap))
    return rc, out.raw[:cap] if rc == 0 else b""


def tv_decode(wire):
    dt, rk, off, nb = C.c_int32(), C.c_int32(), C.c_int64(), C.c_int64()
    dims = (C.c_int64 * 25)()
    rc = lib().orc_tv_decode(wire, C.c_int64(len(wire)), C.byref(dt), C.byref(rk), dims, C.byref(off), C.byref(nb))
    if rc:
        return rc, None
    return 0, (dt.value, [dims[i] for i in range(rk.value)], wire[off.value: off.value + nb.value])


def string_tensor_encode(items):
    n = len(items)
    arr = (C.c_char_p * max(1, n))(*items)
    lens = (C.c_int64 * max(1, n))(*[len(i) for i in items])
    need = lib().orc_string_tensor_encode(C.c_int64(n), arr, lens, None, C.c_int64(0))
    out = C.create_string_buffer(max(int(need), 1))
    got = lib().orc_string_tensor_encode(C.c_int64(n), arr, lens, out, C.c_int64(need))
    return out.raw[:got]


def parse_example_dense_f32(ex, key, has_default=False, default=0.0):
    out = C.c_float()
    rc = lib().orc_parse_example_dense_f32(ex, C.c_int64(len(ex)), key.encode(), int(has_default), C.c_float(default),
                                           C.byref(out))
    return rc, out.value


def regress(string_tensor, n, key, a, b, has_default=False, default=0.0):
    y = np.empty(n, np.float32)
    rc = lib().orc_regress(string_tensor, C.c_int64(len(string_tensor)), C.c_int64(n), key.encode(), int(has_default),
                           C.c_float(default), C.c_float(a), C.c_float(b), _p(y))
    return rc, y.reshape(n, 1)


def tensor_name(s):
    op = C.create_string_buffer(len(s.encode()) + 2)
    idx = C.c_int()
    rc = lib().orc_tensor_name(s.encode(), op, len(op), C.byref(idx))
    return rc, op.value.decode(), idx.value


def jpeg_info(data):
    """(height, width, components) of a baseline JPEG file, or raises ValueError(status)."""
    h, w, c = C.c_int32(), C.c_int32(), C.c_int32()
    rc = lib().orc_jpeg_info(data, C.c_int64(len(data)), C.byref(h), C.byref(w), C.byref(c))
    if rc:
        raise ValueError(rc)
    return h.value, w.value, c.value


def jpeg_decode(data):
    """DecodeJpeg(channels=3): file bytes -> uint8 [H,W,3]."""
    h, w, _ = jpeg_info(data)
    out = np.empty((h, w, 3), np.uint8)
    rc = lib().orc_jpeg_decode(data, C.c_int64(len(data)), _p(out))
    if rc:
        raise ValueError(rc)
    return out
