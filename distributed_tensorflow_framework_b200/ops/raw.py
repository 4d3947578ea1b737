"""Raw (non-autograd) Python wrappers over the C ABI: allocate outputs with torch, launch on the current stream.

Layouts: activations NHWC bf16 (C % 8 == 0); conv filters as 2-D row-major bf16 matrices
[KH*KW*Cin, Cout] (= TF HWIO, /root/reference/distribute_tools.py:71-72) for fprop and
[KH*KW*Cout, Cin] (per-tap transpose) for dgrad / deconv.
"""
import torch

from . import _lib

bf16 = torch.bfloat16
USE_HALO = True  # stride-1 KxK convolutions with C % 64 == 0 use the halo-tile kernel (csrc/conv_halo.cu)


def same_pad(in_size, k, stride):
    """TF 'SAME': out = ceil(in/stride); extra padding goes to the bottom/right. Returns (out, pad_before)."""
    out = -(-in_size // stride)
    total = max((out - 1) * stride + k - in_size, 0)
    return out, total // 2


def _sp():
    return _lib.stream_ptr()


def gemm(a, b, bias=None, relu=False, out_fp32=False, a_transposed=False, out=None, accumulate=False, addend=None,
         stats=None):
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
    if addend is not None:  # act(A @ B + bias + addend), addend bf16 [M, N] laid out like `out`
        assert addend.dtype == bf16 and addend.shape == out.shape and addend.stride(0) == out.stride(0)
    part = None
    if stats is not None:  # dict: receives "part" = [rows, 2, N] per-tile column sums / sums of squares of the output
        rows = 0
        if not (relu or out_fp32 or accumulate or a_transposed or addend is not None):
            rows = _lib.lib().dtf_gemm_stats_rows(a.data_ptr(), b.data_ptr(), out.data_ptr(), M, N, K, a.stride(0),
                                                  b.stride(0), out.stride(0))
        if rows > 0:
            part = torch.empty((rows, 2, N), device=a.device, dtype=torch.float32)
            stats["part"] = part
    _lib.call("dtf_gemm_bf16_ex", a, b, out, bias, addend, part, M, N, K, a.stride(0), b.stride(0), out.stride(0), flags,
              _sp())
    return out


def gemm_ag(a, ag, transposed_shadow, bias=None, relu=False, out_fp32=False):
    """K-AG-GEMM: act(A @ B + bias) with B pulled tile by tile from the owners' heaps (see dtf_gemm_bf16_ag).
    `ag`: dict(world, bounds, off, peers=[ptr per rank], peers_t=[...], shape=(in, out)) attached to an FC variable."""
    import ctypes
    M, K = a.shape
    fin, fout = ag["shape"]
    if transposed_shadow:  # B = W^T [out, in]
        assert K == fout
        N, ptrs = fin, ag["peers_t"]
    else:
        assert K == fin
        N, ptrs = fout, ag["peers"]
    out = torch.empty((M, N), device=a.device, dtype=torch.float32 if out_fp32 else bf16)
    arr = (ctypes.c_void_p * ag["world"])(*ptrs)
    bounds = (ctypes.c_longlong * 9)(*(list(ag["bounds"]) + [ag["bounds"][-1]] * (9 - len(ag["bounds"]))))
    flags = (1 if relu else 0) | (2 if out_fp32 else 0)
    _lib.call("dtf_gemm_bf16_ag", a, arr, out, bias, M, N, K, a.stride(0), N, out.stride(0), flags, ag["world"],
              bounds, ag["off"], 1 if transposed_shadow else 0, _sp())
    return out


def subsample(x, s):
    n, H, W, C = x.shape
    OH, OW = -(-H // s), -(-W // s)
    y = torch.empty((n, OH, OW, C), device=x.device, dtype=bf16)
    _lib.call("dtf_subsample", x, y, n, H, W, C, OH, OW, s, _sp())
    return y


def upsample_zero(dy, in_hw, s):
    n, OH, OW, C = dy.shape
    dx = torch.empty((n, in_hw[0], in_hw[1], C), device=dy.device, dtype=bf16)
    _lib.call("dtf_upsample_zero", dy, dx, n, in_hw[0], in_hw[1], C, OH, OW, s, _sp())
    return dx


# ------------------------------------------------------------------------------------------------ window packing
USE_PACK = True  # strided KxK and few-channel convolutions run as stride-1 halo convolutions over packed windows
USE_TAP_MASKS = True  # skip the structurally zero (tap, phase) weight tiles of space-to-depth packed filters
_PLANS = {}


class _PackPlan(object):
    """Geometry of `P[n, j, i, (r*kwn + q)*cs + c] = x[n, rp*j + r, sx*i + q - ox, c]` and of the stride-1
    convolution over P that equals the original one (see csrc/ops.cu pack_window_kernel)."""

    def __init__(self, H, W, Cx, C, KH, KW, stride, pt, pl, mode):
        self.H, self.W, self.Cx, self.C, self.KH, self.KW, self.stride = H, W, Cx, C, KH, KW, stride
        self.mode = mode
        self.Hp, self.Wq = -(-H // stride), -(-W // stride)
        if mode == "s2d":
            self.rp, self.kwn, self.sx, self.cs, self.ox, self.Cp = 2, 2, 2, Cx, 0, 4 * Cx
        else:
            self.rp, self.kwn, self.sx, self.cs, self.ox, self.Cp = stride, KW, stride, C, pl, 64
        self.taps = []  # (kh, kw, a, b, slot)
        for kh in range(KH):
            a, r = divmod(kh - pt, self.rp)
            for kw in range(KW):
                if mode == "s2d":
                    b, q = divmod(kw - pl, 2)
                else:
                    b, q = 0, kw
                self.taps.append((kh, kw, a, b, r * self.kwn + q))
        amin = min(t[2] for t in self.taps)
        bmin = min(t[3] for t in self.taps)
        self.KH2 = max(t[2] for t in self.taps) - amin + 1
        self.KW2 = max(t[3] for t in self.taps) - bmin + 1
        self.pad_h2, self.pad_w2 = -amin, -bmin
        self.taps = [(kh, kw, (a - amin) * self.KW2 + (b - bmin), slot) for (kh, kw, a, b, slot) in self.taps]
        # space-to-depth: which packed taps carry weights for each of the 4 pixel phases (the halo kernel skips the
        # structurally zero (tap, phase) weight tiles)
        self.n_phase, self.phase_mask = 0, None
        if mode == "s2d":
            import ctypes
            masks = [0, 0, 0, 0]
            for (_, _, tap2, slot) in self.taps:
                masks[slot] |= 1 << tap2
            self.n_phase, self.phase_mask = 4, (ctypes.c_ulonglong * 4)(*masks)
        self._descs = {}

    def descs(self, kind, n_out, device, flip=False):
        """Device block list for dtf_block_copy. kind 'w': [KH*KW*Cx, n_out] -> packed [KH2*KW2*Cp, n_out];
        'wt': per-tap transposed [KH*KW*n_out, Cx] -> [KH2*KW2*n_out, Cp]; 'dw': packed gradient -> filter gradient."""
        key = (kind, n_out, str(device), flip)
        if key not in self._descs:
            rows = []
            for (kh, kw, tap2, slot) in self.taps:
                tsrc = ((self.KH - 1 - kh) * self.KW + (self.KW - 1 - kw)) if flip else (kh * self.KW + kw)
                if kind == "wt":
                    src = tsrc * n_out * self.Cx
                    dst = tap2 * n_out * self.Cp + slot * self.cs
                    r, c = n_out, self.C
                else:
                    src = tsrc * self.Cx * n_out
                    dst = (tap2 * self.Cp + slot * self.cs) * n_out
                    r, c = self.C, n_out
                    if kind == "dw":
                        src, dst = dst, src
                rows.append([src, dst, r | (c << 32)])
            self._descs[key] = (torch.tensor(rows, dtype=torch.int64, device=device), len(rows), rows[0][2] & 0xffffffff,
                                rows[0][2] >> 32)
        return self._descs[key]


def _pack_plan(H, W, Cx, c_used, KH, KW, stride, pads=None):
    """The packing plan for a convolution over x[n, H, W, Cx] (first c_used channels live), or None."""
    if not USE_PACK or KH * KW < 2:
        return None
    if pads is None:
        pads = (same_pad(H, KH, stride)[1], same_pad(W, KW, stride)[1])
    key = (H, W, Cx, c_used, KH, KW, stride, pads)
    if key not in _PLANS:
        if stride == 2 and c_used == Cx and Cx % 16 == 0 and KH > 1 and KW > 1:
            _PLANS[key] = _PackPlan(H, W, Cx, Cx, KH, KW, stride, pads[0], pads[1], "s2d")
        elif stride in (1, 2) and stride * KW * c_used <= 64 and Cx % 64 != 0:
            _PLANS[key] = _PackPlan(H, W, Cx, c_used, KH, KW, stride, pads[0], pads[1], "full")
        else:
            _PLANS[key] = None
    return _PLANS[key]


def pack_window(x, plan):
    n = x.shape[0]
    P = torch.empty((n, plan.Hp, plan.Wq, plan.Cp), device=x.device, dtype=bf16)
    _lib.call("dtf_pack_window", x, P, n, plan.H, plan.W, plan.Cx, plan.C, plan.Hp, plan.Wq, plan.Cp, plan.rp, plan.kwn,
              plan.sx, plan.cs, 0, plan.ox, _sp())
    return P


def _block_copy(src, dst, desc, fp32=False, accumulate=False):
    t, nblk, r, c = desc
    _lib.call("dtf_block_copy", src, src.stride(0), dst, dst.stride(0), t, nblk, r * c, 1 if fp32 else 0,
              1 if accumulate else 0, _sp())


def _packed_fprop(x, w2d, bias, plan, relu, flip=False, keep=None, stats=None):
    """The convolution of `plan` as pack + stride-1 halo convolution; None if the halo kernel does not take it."""
    n_out = w2d.shape[1]
    if not _lib.lib().dtf_conv_halo_supported(plan.Hp, plan.Wq, plan.Cp, plan.KH2, plan.KW2, n_out):
        return None
    n = x.shape[0]
    P = pack_window(x, plan)
    if keep is not None:
        keep["packed"] = P  # the caller may hand it back to conv_wgrad instead of x
    Wp = torch.zeros((plan.KH2 * plan.KW2 * plan.Cp, n_out), device=x.device, dtype=bf16)
    _block_copy(w2d, Wp, plan.descs("w", n_out, x.device, flip))
    out = torch.empty((n, plan.Hp, plan.Wq, n_out), device=x.device, dtype=bf16)
    part = _halo_stats_buffer(stats, relu, n, plan.Hp, plan.Wq, plan.Cp, plan.KH2, plan.KW2, n_out, x.device)
    _lib.call("dtf_conv_halo_bf16_ex", P, Wp, out, bias, part, n, plan.Hp, plan.Wq, plan.Cp, plan.KH2, plan.KW2, n_out,
              plan.pad_h2, plan.pad_w2, 0, 1 if relu else 0, plan.n_phase if USE_TAP_MASKS else 0, plan.phase_mask,
              _sp())
    return out


def _halo_stats_buffer(stats, relu, n, H, W, C, KH, KW, n_out, device):
    """The [rows, 2, n_out] statistics buffer of a halo convolution (registered in `stats`), or None."""
    if stats is None or relu:
        return None
    rows = _lib.lib().dtf_conv_halo_stats_rows(n, H, W, C, KH, KW, n_out)
    if rows <= 0:
        return None
    stats["part"] = torch.empty((rows, 2, n_out), device=device, dtype=torch.float32)
    return stats["part"]


def conv_fprop(x, w2d, bias, KH, KW, stride=1, relu=False, out_fp32=False, c_used=None, keep=None, stats=None):
    """`stats` (dict, optional): where the kernel can, it receives "part" = [rows, 2, n_out] per-tile column sums and
    sums of squares of the output (batch-norm statistics without another pass); absent if the path cannot."""
    n, H, W, C = x.shape
    n_out = w2d.shape[1]
    if KH == 1 and KW == 1 and stride > 1:  # strided 1x1: select the sampled pixels, then a dense GEMM
        return conv_fprop(subsample(x, stride), w2d, bias, 1, 1, 1, relu, out_fp32, stats=stats)
    if KH == 1 and KW == 1 and stride == 1:  # a 1x1 convolution is a dense GEMM: TMA-fed A operand, no gather
        return gemm(x.view(-1, C), w2d, bias, relu=relu, out_fp32=out_fp32, stats=stats).view(n, H, W, n_out)
    OH, pt = same_pad(H, KH, stride)
    OW, pl = same_pad(W, KW, stride)
    if USE_HALO and not out_fp32 and (stride > 1 or C % 64 != 0):
        plan = _pack_plan(H, W, C, C if c_used is None else c_used, KH, KW, stride)
        if plan is not None:
            y = _packed_fprop(x, w2d, bias, plan, relu, keep=keep, stats=stats)
            if y is not None:
                return y
    out = torch.empty((n, OH, OW, n_out), device=x.device, dtype=torch.float32 if out_fp32 else bf16)
    if stride == 1 and not out_fp32 and USE_HALO and _lib.lib().dtf_conv_halo_supported(H, W, C, KH, KW, n_out):
        # halo-tile kernel: one TMA load per tile, all taps are shifted views of the same shared-memory tile
        part = _halo_stats_buffer(stats, relu, n, H, W, C, KH, KW, n_out, x.device)
        _lib.call("dtf_conv_halo_bf16_ex", x, w2d, out, bias, part, n, H, W, C, KH, KW, n_out, pt, pl, 0,
                  1 if relu else 0, 0, None, _sp())
        return out
    flags = (1 if relu else 0) | (2 if out_fp32 else 0)
    _lib.call("dtf_conv_igemm_bf16", x, w2d, out, bias, n, H, W, C, OH, OW, KH, KW, n_out, stride, pt, pl, 0, flags,
              _sp())
    return out


def conv_dgrad(dy, w2d_t, in_hw, KH, KW, stride=1, bias=None, relu=False, addend=None):
    """Data gradient of the SAME conv (H,W)->dy spatial; also the forward of conv2d_transpose. w2d_t: [KH*KW*Cout, Cin].
    `addend` [n, H, W, Cin] is summed into the result: inside the GEMM epilogue for 1x1 layers, by the add kernel else."""
    n, OH, OW, Co = dy.shape
    H, W = in_hw
    Ci = w2d_t.shape[1]
    if addend is not None:
        if KH == 1 and KW == 1 and stride == 1:
            return gemm(dy.view(-1, Co), w2d_t, bias, relu=relu, addend=addend.reshape(-1, Ci)).view(n, H, W, Ci)
        return add(conv_dgrad(dy, w2d_t, in_hw, KH, KW, stride, bias, relu), addend.contiguous())
    if KH == 1 and KW == 1 and stride == 1:
        return gemm(dy.view(-1, Co), w2d_t, bias, relu=relu).view(n, H, W, Ci)
    if KH == 1 and KW == 1 and stride > 1 and bias is None and not relu:
        small = gemm(dy.view(-1, Co), w2d_t).view(n, OH, OW, Ci)
        return upsample_zero(small, (H, W), stride)
    _, pt = same_pad(H, KH, stride)
    _, pl = same_pad(W, KW, stride)
    if USE_HALO and bias is None and not relu:
        if stride == 2:  # space-to-depth: dP = stride-1 data gradient over packed channels, then unpack
            plan = _pack_plan(H, W, Ci, Ci, KH, KW, 2)
            if plan is not None and plan.mode == "s2d" and \
                    _lib.lib().dtf_conv_halo_supported(OH, OW, Co, plan.KH2, plan.KW2, plan.Cp):
                WpT = torch.zeros((plan.KH2 * plan.KW2 * Co, plan.Cp), device=dy.device, dtype=bf16)
                _block_copy(w2d_t, WpT, plan.descs("wt", Co, dy.device))
                dP = torch.empty((n, OH, OW, plan.Cp), device=dy.device, dtype=bf16)
                _lib.call("dtf_conv_halo_bf16_ex", dy, WpT, dP, None, None, n, OH, OW, Co, plan.KH2, plan.KW2, plan.Cp,
                          plan.pad_h2, plan.pad_w2, 1, 0, plan.n_phase if USE_TAP_MASKS else 0, plan.phase_mask, _sp())
                dx = torch.empty((n, H, W, Ci), device=dy.device, dtype=bf16)
                _lib.call("dtf_unpack_s2d", dP, dx, n, H, W, Ci, OH, OW, 0, 0, _sp())
                return dx
        elif stride == 1 and Co % 64 != 0:  # few output channels: the gradient is a convolution of dy with flipped taps
            plan = _pack_plan(H, W, Co, Co, KH, KW, 1, pads=(KH - 1 - pt, KW - 1 - pl))
            if plan is not None:
                dx = _packed_fprop(dy, w2d_t, None, plan, False, flip=True)
                if dx is not None:
                    return dx
    dx = torch.empty((n, H, W, Ci), device=dy.device, dtype=bf16)
    if stride == 1 and USE_HALO and _lib.lib().dtf_conv_halo_supported(H, W, Co, KH, KW, Ci):
        _lib.call("dtf_conv_halo_bf16", dy, w2d_t, dx, bias, n, H, W, Co, KH, KW, Ci, pt, pl, 1, 1 if relu else 0, _sp())
        return dx
    _lib.call("dtf_conv_igemm_bf16", dy, w2d_t, dx, bias, n, OH, OW, Co, H, W, KH, KW, Ci, stride, pt, pl, 1,
              1 if relu else 0, _sp())
    return dx


def conv_wgrad(x, dy, KH, KW, stride=1, out=None, c_used=None, packed=None, x_shape=None):
    """dW[KH*KW*Cin, Cout] fp32 accumulated into `out` (zeroed if not given). `packed` (with `x_shape`) is the packed
    window tensor conv_fprop produced for the same x; x itself may then be None."""
    n, H, W, C = x.shape if x is not None else x_shape
    _, OH, OW, Co = dy.shape
    _, pt = same_pad(H, KH, stride)
    _, pl = same_pad(W, KW, stride)
    if out is None:
        out = torch.zeros((KH * KW * C, Co), device=dy.device, dtype=torch.float32)
    if KH == 1 and KW == 1 and stride > 1:
        x = subsample(x, stride)
        stride = 1
    if KH == 1 and KW == 1 and stride == 1:  # dW = x^T dy: dense MN-major A operand, split over pixels
        return gemm(x.reshape(-1, C), dy.view(-1, Co), a_transposed=True, out=out, accumulate=True)
    if stride == 1 and USE_HALO and _lib.lib().dtf_conv_wgrad_halo_supported(H, W, C, KH, KW, Co):
        _lib.call("dtf_conv_wgrad_halo_bf16", x, dy, out, n, H, W, C, KH, KW, Co, pt, pl, _sp())
        return out
    if (USE_HALO or packed is not None) and (stride > 1 or C % 64 != 0):
        plan = _pack_plan(H, W, C, C if c_used is None else c_used, KH, KW, stride)
        halo_ok = plan is not None and _lib.lib().dtf_conv_wgrad_halo_supported(plan.Hp, plan.Wq, plan.Cp, plan.KH2,
                                                                                plan.KW2, Co)
        if plan is not None and (halo_ok or packed is not None):
            P = pack_window(x, plan) if packed is None else packed
            dWp = torch.zeros((plan.KH2 * plan.KW2 * plan.Cp, Co), device=dy.device, dtype=torch.float32)
            if halo_ok:
                _lib.call("dtf_conv_wgrad_halo_bf16", P, dy, dWp, n, plan.Hp, plan.Wq, plan.Cp, plan.KH2, plan.KW2, Co,
                          plan.pad_h2, plan.pad_w2, _sp())
            else:  # small maps: the gather kernel on the packed (stride-1) problem
                _lib.call("dtf_conv_wgrad_bf16", P, dy, dWp, n, plan.Hp, plan.Wq, plan.Cp, plan.Hp, plan.Wq, plan.KH2,
                          plan.KW2, Co, 1, plan.pad_h2, plan.pad_w2, _sp())
            _block_copy(dWp, out, plan.descs("dw", Co, dy.device), fp32=True, accumulate=True)
            return out
    assert x is not None, "conv_wgrad: x is required when no packed tensor fits the geometry"
    _lib.call("dtf_conv_wgrad_bf16", x, dy, out, n, H, W, C, OH, OW, KH, KW, Co, stride, pt, pl, _sp())
    return out


# ------------------------------------------------------------------------------------------------ memory-bound ops
def convert_input(src, c_pad=None, scale=1.0, shift=0.0):
    """[..., C] fp32/uint8 -> [..., Cpad] bf16 (normalise + pad channels to a multiple of 8)."""
    C = src.shape[-1]
    if c_pad is None:
        c_pad = (C + 7) // 8 * 8
    src = src.contiguous()
    out = torch.empty(src.shape[:-1] + (c_pad,), device=src.device, dtype=bf16)
    _lib.call("dtf_convert_input", src, out, src.numel() // C, C, c_pad, scale, shift,
              1 if src.dtype == torch.uint8 else 0, _sp())
    return out


def relu(x):
    y = torch.empty_like(x)
    _lib.call("dtf_relu_fwd", x, y, x.numel(), _sp())
    return y


def relu_bwd(dy, y):
    dx = torch.empty_like(dy)
    _lib.call("dtf_relu_bwd", dy, y, dx, dy.numel(), _sp())
    return dx


def add(a, b, relu=False):
    z = torch.empty_like(a)
    _lib.call("dtf_add", a, b, z, a.numel(), 1 if relu else 0, _sp())
    return z


def bias_grad(dy2d, out=None):
    R, N = dy2d.shape
    if out is None:
        out = torch.zeros(N, device=dy2d.device, dtype=torch.float32)
    _lib.call("dtf_bias_grad", dy2d, out, R, N, _sp())
    return out


def relu_bwd_bias_grad(dy2d, y2d, db=None):
    R, N = dy2d.shape
    if db is None:
        db = torch.zeros(N, device=dy2d.device, dtype=torch.float32)
    dz = torch.empty_like(dy2d)
    _lib.call("dtf_relu_bwd_bias_grad", dy2d, y2d, dz, db, R, N, _sp())
    return dz, db


def pool_geometry(H, W, k, s):
    OH, pt = same_pad(H, k[0], s[0])
    OW, pl = same_pad(W, k[1], s[1])
    return OH, OW, pt, pl


def pool_fwd(x, k=(2, 2), s=(2, 2), is_max=True, want_idx=True, pre_scale=None, pre_shift=None):
    """Returns (y, idx); idx (uint8 argmax positions) is None for average pooling. With pre_scale / pre_shift the
    pooled quantity is relu(x * scale + shift) (batch norm + ReLU applied on the fly)."""
    n, H, W, C = x.shape
    OH, OW, pt, pl = pool_geometry(H, W, k, s)
    y = torch.empty((n, OH, OW, C), device=x.device, dtype=bf16)
    idx = torch.empty((n, OH, OW, C), device=x.device, dtype=torch.uint8) if (is_max and want_idx) else None
    _lib.call("dtf_pool_fwd_affine", x, y, idx, pre_scale, pre_shift, n, H, W, C, OH, OW, k[0], k[1], s[0], s[1], pt, pl,
              1 if is_max else 0, _sp())
    return y, idx


def pool_bwd(idx, dy, in_shape, k=(2, 2), s=(2, 2), is_max=True):
    n, H, W, C = in_shape
    OH, OW, pt, pl = pool_geometry(H, W, k, s)
    dx = torch.empty(tuple(in_shape), device=dy.device, dtype=bf16)
    _lib.call("dtf_pool_bwd", idx, dy, dx, n, H, W, C, OH, OW, k[0], k[1], s[0], s[1], pt, pl, 1 if is_max else 0, _sp())
    return dx


def lrn_fwd(x, radius=4, bias=1.0, alpha=0.001 / 9.0, beta=0.75):
    y = torch.empty_like(x)
    C = x.shape[-1]
    _lib.call("dtf_lrn_fwd", x, y, x.numel() // C, C, radius, bias, alpha, beta, _sp())
    return y


def lrn_bwd(x, dy, radius=4, bias=1.0, alpha=0.001 / 9.0, beta=0.75):
    dx = torch.empty_like(x)
    C = x.shape[-1]
    _lib.call("dtf_lrn_bwd", x, dy, dx, x.numel() // C, C, radius, bias, alpha, beta, _sp())
    return dx


def _part(R, N, device):
    nblk = _lib.lib().dtf_colreduce_blocks(R, N)
    return torch.empty(nblk, 2, N, device=device, dtype=torch.float32)


def bn_train_fwd(x2d, gamma, beta, run_mean, run_var, eps=1e-3, momentum=0.1, res=None, relu=False, part=None,
                 want_mask=False):
    """Training-mode batch norm over rows of x2d [R, N]. Returns (y, mean, rstd, scale, shift). `part` [rows, 2, N]:
    per-tile sums / sums of squares of x2d already produced by the kernel that wrote it (no statistics pass then)."""
    R, N = x2d.shape
    dev = x2d.device
    stat = torch.empty(4, N, device=dev, dtype=torch.float32)
    mean, rstd, scale, shift = stat[0], stat[1], stat[2], stat[3]
    if part is not None:
        assert part.shape[1:] == (2, N) and part.dtype == torch.float32
        _lib.call("dtf_bn_finalize_rows", part, part.shape[0], gamma, beta, mean, rstd, scale, shift, run_mean, run_var,
                  N, R, eps, momentum, _sp())
    else:
        part = _part(R, N, dev)
        _lib.call("dtf_bn_stats", x2d, part, R, N, _sp())
        _lib.call("dtf_bn_finalize", part, gamma, beta, mean, rstd, scale, shift, run_mean, run_var, N, R, eps,
                  momentum, _sp())
    y = torch.empty_like(x2d)
    if want_mask and relu:
        # the ReLU decisions as one bit per element: what the backward needs of y when a residual went into it
        mask = torch.empty((R, N // 8), device=dev, dtype=torch.uint8)
        _lib.call("dtf_bn_apply_mask", x2d, res, y, mask, scale, shift, R, N, 1, _sp())
        return y, mean, rstd, scale, shift, mask
    _lib.call("dtf_bn_apply", x2d, res, y, scale, shift, R, N, 1 if relu else 0, _sp())
    return y, mean, rstd, scale, shift


def bn_stats_coeffs(x2d, gamma, beta, run_mean, run_var, eps=1e-3, momentum=0.1, part=None):
    """Training-mode statistics of x2d [R, N] folded into (mean, rstd, scale, shift); updates the running statistics.
    `part`: per-tile partial sums already produced by the kernel that wrote x2d (no statistics pass then)."""
    R, N = x2d.shape
    stat = torch.empty(4, N, device=x2d.device, dtype=torch.float32)
    mean, rstd, scale, shift = stat[0], stat[1], stat[2], stat[3]
    if part is not None and tuple(part.shape[1:]) == (2, N):
        _lib.call("dtf_bn_finalize_rows", part, part.shape[0], gamma, beta, mean, rstd, scale, shift, run_mean, run_var,
                  N, R, eps, momentum, _sp())
    else:
        part = _part(R, N, x2d.device)
        _lib.call("dtf_bn_stats", x2d, part, R, N, _sp())
        _lib.call("dtf_bn_finalize", part, gamma, beta, mean, rstd, scale, shift, run_mean, run_var, N, R, eps,
                  momentum, _sp())
    return mean, rstd, scale, shift


def bn_apply_dual(x2d, res2d, scale, shift, res_scale, res_shift, relu=True):
    """y = act(x*scale + shift + res*res_scale + res_shift) (+ the ReLU bitmask): two batch norms joined by a residual
    add in one pass; the normalised residual branch is never written."""
    R, N = x2d.shape
    y = torch.empty_like(x2d)
    mask = torch.empty((R, N // 8), device=x2d.device, dtype=torch.uint8) if relu else None
    _lib.call("dtf_bn_apply_dual", x2d, res2d, y, mask, scale, shift, res_scale, res_shift, R, N, 1 if relu else 0,
              _sp())
    return y, mask


def bn_eval_fwd(x2d, gamma, beta, run_mean, run_var, eps=1e-3, res=None, relu=False):
    R, N = x2d.shape
    scale = torch.empty(N, device=x2d.device, dtype=torch.float32)
    shift = torch.empty_like(scale)
    _lib.call("dtf_bn_fold", run_mean, run_var, gamma, beta, scale, shift, N, eps, _sp())
    y = torch.empty_like(x2d)
    _lib.call("dtf_bn_apply", x2d, res, y, scale, shift, R, N, 1 if relu else 0, _sp())
    return y


def bn_bwd(dy2d, y2d, x2d, mean, rstd, gamma, relu=False, dgamma_acc=None, dbeta_acc=None, mask_scale=None,
           mask_shift=None):
    """Returns (dx, dz, dgamma, dbeta); dz = masked upstream gradient (also the residual-branch gradient).
    The parameter gradients are ADDED into dgamma_acc / dbeta_acc (pass gradient-arena views to avoid extra passes).
    With relu and mask_scale / mask_shift (the forward's folded coefficients, no residual) the ReLU mask is
    recomputed from x: y is not read and dz is not materialised (returned as None)."""
    R, N = x2d.shape
    dev = x2d.device
    if dgamma_acc is None:
        acc = torch.zeros(2, N, device=dev, dtype=torch.float32)
        dgamma_acc, dbeta_acc = acc[0], acc[1]
    part = _part(R, N, dev)
    from_x = relu and mask_scale is not None
    mode = 2 if from_x else (1 if relu else 0)
    if mode == 1 and y2d is not None and y2d.dtype == torch.uint8:
        mode = 3  # y2d is the forward's ReLU bitmask [R, N/8]
    dz = None if from_x else (torch.empty_like(dy2d) if relu else dy2d)
    _lib.call("dtf_bn_bwd_reduce", dy2d, y2d if mode in (1, 3) else None, x2d, dz if mode in (1, 3) else None, part,
              mean, rstd,
              mask_scale if from_x else None, mask_shift if from_x else None, R, N, mode, _sp())
    dx = torch.empty_like(x2d)
    coef = torch.empty(3, N, device=dev, dtype=torch.float32)
    _lib.call("dtf_bn_bwd_apply", dy2d if from_x else dz, x2d, dx, part, mean, rstd, gamma, coef, dgamma_acc, dbeta_acc,
              mask_scale if from_x else None, mask_shift if from_x else None, R, N, _sp())
    return dx, dz, dgamma_acc, dbeta_acc


_scratch = {}


def loss_scratch(device, nfloats):
    """Per-device zero-initialised scratch for the loss kernels (word 0 is the arrival counter)."""
    key = (device.index if device.index is not None else torch.cuda.current_device())
    buf = _scratch.get(key)
    if buf is None or buf.numel() < nfloats:
        buf = torch.zeros(max(nfloats, 1 + _lib.lib().dtf_loss_scratch_blocks(), 65537), device=device,
                          dtype=torch.float32)
        _scratch[key] = buf
    return buf


def softmax_xent(logits, labels, gscale=1.0, want_grad=True):
    """Fused mean softmax cross-entropy + gradient. logits [B,C] bf16/fp32; labels int64 [B]."""
    B, C = logits.shape
    logits = logits.contiguous()
    loss = torch.empty(1, device=logits.device, dtype=torch.float32)
    dl = torch.empty_like(logits) if want_grad else None
    _lib.call("dtf_softmax_xent", logits, labels, dl, loss, loss_scratch(logits.device, 1 + B), B, C, gscale,
              1 if logits.dtype == bf16 else 0, _sp())
    return loss, dl


def mse(pred, target, gscale=1.0, want_grad=True):
    pred = pred.contiguous()
    target = target.contiguous()
    loss = torch.empty(1, device=pred.device, dtype=torch.float32)
    dp = torch.empty_like(pred) if want_grad else None
    _lib.call("dtf_mse", pred, target, dp, loss, loss_scratch(pred.device, 0), pred.numel(), gscale,
              1 if pred.dtype == bf16 else 0, _sp())
    return loss, dp


def l2_loss(w, scale=1.0):
    out = torch.empty(1, device=w.device, dtype=torch.float32)
    _lib.call("dtf_l2_loss", w, out, loss_scratch(w.device, 0), w.numel(), scale, _sp())
    return out


def top1_hits(logits, labels):
    hits = torch.zeros(1, device=logits.device, dtype=torch.int32)
    B, C = logits.shape
    _lib.call("dtf_top1_hits", logits.contiguous(), labels, hits, B, C, 1 if logits.dtype == bf16 else 0, _sp())
    return hits


OPTIM_KINDS = {"sgd": 0, "momentum": 1, "adam": 2}


def optim_step(kind, p, g, m, v, shadow, lr, b1=0.9, b2=0.999, eps=1e-8, wd=0.0, gscale=1.0, step=1, hyper=None):
    _lib.call("dtf_optim_step", OPTIM_KINDS[kind], p, g, m, v, shadow, p.numel(), lr, b1, b2, eps, wd, gscale, step,
              hyper, _sp())


def init_(t, kind, a=0.0, b=1.0, seed=0, stream_id=0):
    """kind: 'constant' | 'uniform' | 'normal' | 'truncated_normal' (fp32 tensor, in place)."""
    k = {"constant": 0, "uniform": 1, "normal": 2, "truncated_normal": 3}[kind]
    _lib.call("dtf_init", t, t.numel(), k, a, b, seed, stream_id, _sp())
    return t
