"""Raw (non-autograd) Python wrappers over the C ABI: allocate outputs with torch, launch on the current stream.

Layouts: activations NHWC bf16 (C % 8 == 0); conv filters as 2-D row-major bf16 matrices
[KH*KW*Cin, Cout] (= TF HWIO, /root/reference/distribute_tools.py:71-72) for fprop and
[KH*KW*Cout, Cin] (per-tap transpose) for dgrad / deconv.
"""
import torch

from . import _lib

bf16 = torch.bfloat16


def same_pad(in_size, k, stride):
    """TF 'SAME': out = ceil(in/stride); extra padding goes to the bottom/right. Returns (out, pad_before)."""
    out = -(-in_size // stride)
    total = max((out - 1) * stride + k - in_size, 0)
    return out, total // 2


def _sp():
    return _lib.stream_ptr()


def gemm(a, b, bias=None, relu=False, out_fp32=False, a_transposed=False, out=None, accumulate=False):
    """act(A @ B + bias). a: [M,K] (or [K,M] if a_transposed), b: [K,N], both bf16 row-major."""
    if a_transposed:
        K, M = a.shape
    else:
        M, K = a.shape
    K2, N = b.shape
    assert K == K2, (a.shape, b.shape)
    if out is None:
        out = torch.empty((M, N), device=a.device, dtype=torch.float32 if (out_fp32 or accumulate) else bf16)
    flags = (1 if relu else 0) | (2 if (out_fp32 or accumulate) else 0) | (4 if accumulate else 0) | \
            (8 if a_transposed else 0)
    _lib.call("dtf_gemm_bf16", a, b, out, bias, M, N, K, a.stride(0), b.stride(0), out.stride(0), flags, _sp())
    return out


def conv_fprop(x, w2d, bias, KH, KW, stride=1, relu=False, out_fp32=False):
    n, H, W, C = x.shape
    n_out = w2d.shape[1]
    OH, pt = same_pad(H, KH, stride)
    OW, pl = same_pad(W, KW, stride)
    out = torch.empty((n, OH, OW, n_out), device=x.device, dtype=torch.float32 if out_fp32 else bf16)
    flags = (1 if relu else 0) | (2 if out_fp32 else 0)
    _lib.call("dtf_conv_igemm_bf16", x, w2d, out, bias, n, H, W, C, OH, OW, KH, KW, n_out, stride, pt, pl, 0, flags,
              _sp())
    return out


def conv_dgrad(dy, w2d_t, in_hw, KH, KW, stride=1, bias=None, relu=False):
    """Data gradient of the SAME conv (H,W)->dy spatial; also the forward of conv2d_transpose. w2d_t: [KH*KW*Cout, Cin]."""
    n, OH, OW, Co = dy.shape
    H, W = in_hw
    Ci = w2d_t.shape[1]
    _, pt = same_pad(H, KH, stride)
    _, pl = same_pad(W, KW, stride)
    dx = torch.empty((n, H, W, Ci), device=dy.device, dtype=bf16)
    _lib.call("dtf_conv_igemm_bf16", dy, w2d_t, dx, bias, n, OH, OW, Co, H, W, KH, KW, Ci, stride, pt, pl, 1,
              1 if relu else 0, _sp())
    return dx


def conv_wgrad(x, dy, KH, KW, stride=1, out=None):
    """dW[KH*KW*Cin, Cout] fp32 accumulated into `out` (zeroed if not given)."""
    n, H, W, C = x.shape
    _, OH, OW, Co = dy.shape
    _, pt = same_pad(H, KH, stride)
    _, pl = same_pad(W, KW, stride)
    if out is None:
        out = torch.zeros((KH * KW * C, Co), device=x.device, dtype=torch.float32)
    _lib.call("dtf_conv_wgrad_bf16", x, dy, out, n, H, W, C, OH, OW, KH, KW, Co, stride, pt, pl, _sp())
    return out
