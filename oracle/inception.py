"""Python binding of oracle/inception_oracle.c (test infrastructure only)."""
import ctypes as C

import numpy as np

from . import lib, num_threads


def conv_params(seed, name, cout, kh, kw, cin):
    w = np.empty((cout, kh, kw, cin), np.float32)
    b = np.empty(cout, np.float32)
    lib().orc_inc_conv_params(C.c_uint64(seed), name.encode(), cout, kh, kw, cin, w.ctypes.data_as(C.c_void_p),
                              b.ctypes.data_as(C.c_void_p))
    return w, b


def fc_params(seed, classes, cin=2048):
    w = np.empty((classes, cin), np.float32)
    b = np.empty(classes, np.float32)
    lib().orc_inc_fc_params(C.c_uint64(seed), classes, cin, w.ctypes.data_as(C.c_void_p), b.ctypes.data_as(C.c_void_p))
    return w, b


def conv2d_nhwc(x, w_ohwi, bias, stride=1, pad=(0, 0), relu=True, add_bias=True, round_bf16=False, threads=None):
    x = np.ascontiguousarray(x, np.float32)
    w = np.ascontiguousarray(w_ohwi, np.float32)
    n, h, wd, cin = x.shape
    cout, kh, kw, _ = w.shape
    oh, ow = (h + 2 * pad[0] - kh) // stride + 1, (wd + 2 * pad[1] - kw) // stride + 1
    out = np.empty((n, oh, ow, cout), np.float32)
    b = np.ascontiguousarray(bias, np.float32)
    rc = lib().orc_conv2d_nhwc(x.ctypes.data_as(C.c_void_p), n, h, wd, cin, w.ctypes.data_as(C.c_void_p),
                               b.ctypes.data_as(C.c_void_p), cout, kh, kw, stride, stride, pad[0], pad[1], int(add_bias),
                               int(relu), int(round_bf16), out.ctypes.data_as(C.c_void_p), cout, 0,
                               threads or num_threads())
    if rc:
        raise ValueError(f"oracle conv: status {rc}")
    return out


class Net:
    """Inception-v3 (random-init from seed).  mode 0 = fp32, 1 = bf16 emulation."""

    def __init__(self, classes=1000, seed=2024, mode=0):
        self.classes, self.seed, self.mode = classes, seed, mode

    def forward(self, images, threads=None, tap=None):
        x = np.ascontiguousarray(images, np.float32)
        n, side = x.shape[0], x.shape[1]
        logits = np.empty((n, self.classes), np.float32)
        dims = (C.c_int * 4)()
        th = threads or num_threads()
        args = [C.c_uint64(self.seed), self.classes, self.mode, x.ctypes.data_as(C.c_void_p), n, side,
                logits.ctypes.data_as(C.c_void_p), th]
        if tap is None:
            rc = lib().orc_inc_forward(*args, None, None, C.c_int64(0), dims)
            if rc:
                raise ValueError(f"oracle inception: status {rc}")
            return logits
        rc = lib().orc_inc_forward(*args, tap.encode(), None, C.c_int64(0), dims)     # size query
        if rc not in (0, 11):
            raise KeyError(f"no activation named {tap} (status {rc})")
        shape = tuple(dims[i] for i in range(4))
        out = np.empty(shape, np.float32)
        rc = lib().orc_inc_forward(*args, tap.encode(), out.ctypes.data_as(C.c_void_p), C.c_int64(out.size), dims)
        if rc:
            raise ValueError(f"oracle inception: status {rc}")
        return logits, out
