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
