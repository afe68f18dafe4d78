"""Device-level entry points of the C ABI on torch CUDA tensors (torch only owns the memory and the stream)."""
import ctypes as C

import numpy as np
import torch

from . import _abi
from .host import _check


def _stream():
    return C.c_void_p(torch.cuda.current_stream().cuda_stream)


def _need_cuda(*ts):
    for t in ts:
        if t is not None and (not t.is_cuda or not t.is_contiguous()):
            raise ValueError("device entry points take contiguous CUDA tensors")


def affine_f32(x, a, b, out=None):
    """K1: y = a*x + b (two roundings) on a CUDA float32 tensor."""
    _need_cuda(x, out)
    y = out if out is not None else torch.empty_like(x)
    _check(_abi.load_library().ftf_affine_f32_dev(x.data_ptr(), y.data_ptr(), x.numel(), a, b, _stream()))
    return y


def resize_norm(images_u8, out_h, out_w, mean, scale, out_dtype=torch.float32, out=None):
    """K2: u8 [N,H,W,3] -> [N,out_h,out_w,3] float32/bfloat16."""
    _need_cuda(images_u8, out)
    n, h, w, c = images_u8.shape
    if c != 3 or images_u8.dtype != torch.uint8:
        raise ValueError("images must be uint8 [N,H,W,3]")
    y = out if out is not None else torch.empty((n, out_h, out_w, 3), dtype=out_dtype, device=images_u8.device)
    code = _abi.FLOAT if y.dtype == torch.float32 else _abi.BFLOAT16
    _check(_abi.load_library().ftf_resize_norm_dev(images_u8.data_ptr(), n, h, w, out_h, out_w, mean, scale,
                                                   y.data_ptr(), code, _stream()))
    return y


def softmax_topk(x, k, n_labels=None, apply_softmax=True, want_probs=True):
    """K7 on a CUDA float32 [M,N] tensor -> (probs or None, idx int32 [M,k], score float32 [M,k])."""
    _need_cuda(x)
    m, n = x.shape
    probs = torch.empty_like(x) if (want_probs and apply_softmax) else None
    idx = torch.empty((m, k), dtype=torch.int32, device=x.device)
    score = torch.empty((m, k), dtype=torch.float32, device=x.device)
    _check(_abi.load_library().ftf_softmax_topk_dev(
        x.data_ptr(), m, n, n if n_labels is None else n_labels, k, 1 if apply_softmax else 0,
        probs.data_ptr() if probs is not None else None, idx.data_ptr(), score.data_ptr(), _stream()))
    return probs, idx, score


def top_labels(scores_np, n_labels, k):
    """toTextLabels' selection for a host score matrix: upload, K7 without softmax, download."""
    x = torch.from_numpy(np.ascontiguousarray(scores_np, dtype=np.float32)).cuda()
    _, idx, score = softmax_topk(x, k, n_labels=n_labels, apply_softmax=False)
    torch.cuda.synchronize()
    return idx.cpu().numpy(), score.cpu().numpy()


def conv2d_bf16(x, w, bias, stride=1, pad=(0, 0), relu=True, out_dtype=torch.bfloat16, out=None, out_off=0):
    """K3: x bf16 [N,H,W,Cin], w bf16 [Cout,KH,KW,Cin], bias f32 [Cout] or None -> [N,OH,OW,Cout] (or a slice of out)."""
    _need_cuda(x, w, bias, out)
    n, h, wd, cin = x.shape
    cout, kh, kw, cin2 = w.shape
    if cin != cin2 or x.dtype != torch.bfloat16 or w.dtype != torch.bfloat16:
        raise ValueError("conv2d_bf16: bf16 NHWC input and [Cout,KH,KW,Cin] weights")
    oh, ow = (h + 2 * pad[0] - kh) // stride + 1, (wd + 2 * pad[1] - kw) // stride + 1
    y = out if out is not None else torch.empty((n, oh, ow, cout), dtype=out_dtype, device=x.device)
    code = _abi.FLOAT if y.dtype == torch.float32 else _abi.BFLOAT16
    _check(_abi.load_library().ftf_conv2d_bf16_dev(
        x.data_ptr(), n, h, wd, cin, cin, w.data_ptr(), bias.data_ptr() if bias is not None else None, cout, kh, kw,
        stride, pad[0], pad[1], 1 if relu else 0, y.data_ptr(), code, y.shape[-1], out_off, _stream()))
    return y


def maxpool3x3s2(x, out=None, out_off=0):
    """K4 on bf16 NHWC."""
    _need_cuda(x, out)
    n, h, w, c = x.shape
    y = out if out is not None else torch.empty((n, (h - 3) // 2 + 1, (w - 3) // 2 + 1, c), dtype=torch.bfloat16, device=x.device)
    _check(_abi.load_library().ftf_maxpool3x3s2_bf16_dev(x.data_ptr(), n, h, w, c, y.data_ptr(), y.shape[-1], out_off, _stream()))
    return y


def pool_branch(x, w, bias, out=None, out_off=0):
    """K5: relu(avgpool3x3(conv1x1(x)) + bias); x bf16 [N,H,W,Cin], w bf16 [Cout,Cin], bias f32 [Cout]."""
    _need_cuda(x, w, bias, out)
    n, h, wd, cin = x.shape
    cout = w.shape[0]
    y = out if out is not None else torch.empty((n, h, wd, cout), dtype=torch.bfloat16, device=x.device)
    _check(_abi.load_library().ftf_pool_branch_bf16_dev(x.data_ptr(), n, h, wd, cin, cin, w.data_ptr(), bias.data_ptr(), cout,
                                                        y.data_ptr(), y.shape[-1], out_off, _stream()))
    return y


def global_avgpool(x):
    """K6: bf16 [N,H,W,C] -> bf16 [N,C]."""
    _need_cuda(x)
    n, h, w, c = x.shape
    y = torch.empty((n, c), dtype=torch.bfloat16, device=x.device)
    _check(_abi.load_library().ftf_global_avgpool_bf16_dev(x.data_ptr(), n, h * w, c, y.data_ptr(), _stream()))
    return y


def jpeg_info(data):
    """(height, width, components) from the marker segments of one JPEG file; host only."""
    h, w, c = C.c_int32(), C.c_int32(), C.c_int32()
    buf = np.frombuffer(data, np.uint8)
    _check(_abi.load_library().ftf_jpeg_info(buf.ctypes.data if buf.size else None, buf.size, C.byref(h), C.byref(w), C.byref(c)))
    return h.value, w.value, c.value


class JpegBatch:
    """J1-J3 behind ftf_jpeg_batch_*: a reusable batch of JPEG files decoded on the current CUDA device."""

    def __init__(self, device=None):
        self._lib = _abi.load_library()
        self._h = C.c_void_p()
        dev = torch.cuda.current_device() if device is None else device
        _check(self._lib.ftf_jpeg_batch_create(dev, C.byref(self._h)))
        self.n = 0

    def _ck(self, rc):
        if rc:
            msg = self._lib.ftf_jpeg_batch_error(self._h)
            from .host import _ERRORS
            raise _ERRORS.get(rc, RuntimeError)((msg or b"").decode("utf-8", "replace") + f" [status {rc}]")

    def load(self, files):
        """files: list of bytes.  Parses the headers, stages and uploads the entropy-coded data on the current stream."""
        n = len(files)
        self._keep = [np.frombuffer(f, np.uint8) for f in files]
        ptrs = (C.c_void_p * max(1, n))(*[k.ctypes.data for k in self._keep])
        lens = (C.c_int64 * max(1, n))(*[k.size for k in self._keep])
        offs = (C.c_int64 * max(1, n))()
        hs = (C.c_int32 * max(1, n))()
        ws = (C.c_int32 * max(1, n))()
        self._ck(self._lib.ftf_jpeg_batch_load(self._h, n, ptrs, lens, offs, hs, ws, _stream()))
        self.n = n
        self.offsets = [offs[i] for i in range(n)]
        self.shapes = [(hs[i], ws[i]) for i in range(n)]
        return self

    @property
    def rgb_bytes(self):
        return self._lib.ftf_jpeg_batch_rgb_bytes(self._h)

    @property
    def uploaded_bytes(self):
        return self._lib.ftf_jpeg_batch_scan_bytes(self._h)

    @property
    def segments(self):
        return self._lib.ftf_jpeg_batch_segments(self._h)

    def decode(self, out=None):
        """Enqueues the three kernels on the current stream; returns the flat uint8 output tensor."""
        if out is None:
            out = torch.empty(max(1, self.rgb_bytes), dtype=torch.uint8, device="cuda")
        _need_cuda(out)
        self._ck(self._lib.ftf_jpeg_batch_decode_dev(self._h, out.data_ptr(), out.numel(), _stream()))
        return out

    def status(self):
        """Waits for the decode; per-image status codes.  Raises ValueError for the first corrupt image."""
        st = (C.c_int32 * max(1, self.n))()
        rc = self._lib.ftf_jpeg_batch_status(self._h, st, self.n)
        self.last_status = [st[i] for i in range(self.n)]
        self._ck(rc)
        return self.last_status

    def images(self, flat):
        """Views of the decoded images, uint8 [h,w,3] each, into the flat output tensor."""
        return [flat[o: o + h * w * 3].view(h, w, 3) for o, (h, w) in zip(self.offsets, self.shapes)]

    def close(self):
        if self._h:
            self._lib.ftf_jpeg_batch_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


def decode_jpeg(files):
    """DecodeJpeg(channels=3) for a list of files -> list of CUDA uint8 [h,w,3] tensors."""
    b = JpegBatch()
    try:
        flat = b.load(files).decode()
        b.status()
        return b.images(flat)
    finally:
        b.close()
