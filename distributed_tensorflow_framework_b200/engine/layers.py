"""Layer library ("tools"): same function names, argument order, defaults and variable names as the reference.

Parity: /root/reference/distribute_tools.py:24-224 -- `_variable_on_cpu`, `_variable_with_weight_decay`, `conv`,
`conv_eval`, `conv_nonacti`, `conv_nonacti_eval`, `acti_layer`, `deconv`, `deconv_eval`, `pool`, `batch_norm`,
`norm`, `FC_layer`, `weight`, `bias`. Tensors are NHWC. On a CUDA tensor every op runs a hand-written sm_100a
kernel (engine/functional.py); on a CPU tensor it runs the plain PyTorch fp32 op that also serves as the oracle.

Deliberate fixes: `norm` really applies LRN unless --lrn_compat_noop (Q15); `batch_norm` keeps the
literal axis-0 / no-affine behaviour and `batch_norm_channel` adds the standard per-channel BN with affine and
running statistics (Q16); `deconv` documents that dim 2 of its filter is the OUTPUT channel count (Q17).
Extensions used by the example models: `fc_nonacti`, `global_avg_pool`, `add_relu`, `batch_norm_channel`.
"""
import torch
import torch.nn.functional as F

from ..api import flags as _flags
from . import functional as Fn
from . import variables as V
from .variables import (ConstantInitializer, TruncatedNormalInitializer, XavierInitializer, variable_scope)

_training = True
_emulate_bf16 = False


class _RoundBF16(torch.autograd.Function):
    """Oracle helper: rounds a tensor to bf16 precision in the forward pass (optional) and its gradient in the
    backward pass -- the two places where the CUDA path stores an activation / a data gradient as bf16."""

    @staticmethod
    def forward(ctx, x, fwd):
        return x.to(torch.bfloat16).float() if fwd else x.view_as(x)

    @staticmethod
    def backward(ctx, g):
        return g.to(torch.bfloat16).float(), None


def emulate_bf16(flag: bool):
    """CPU oracle mode: every layer output (and the gradient flowing back into it) is rounded to bf16 exactly where
    the CUDA kernels store bf16, while all arithmetic stays fp32. With identical (bf16-representable) weights the
    CPU path then reproduces the CUDA forward pass almost bit for bit, so ReLU / max-pool decisions agree and one
    training step can be compared per variable with a tight bound (tests/test_gpu_round2.py). Returns the old flag."""
    global _emulate_bf16
    old, _emulate_bf16 = _emulate_bf16, bool(flag)
    return old


def _q(x, fwd=True):
    return _RoundBF16.apply(x, fwd) if _emulate_bf16 else x


def set_training(flag: bool):
    """Train / eval switch for batch-norm statistics (the reference has no inference statistics, Q16)."""
    global _training
    _training = bool(flag)


def is_training():
    return _training


def _store():
    return V.default_store()


def _get(name, shape, initializer, kind="generic", store_shape=None, wd=None, trainable=True):
    st = _store()
    v = st.get_variable(name, shape, initializer, trainable=trainable, kind=kind, store_shape=store_shape, wd=wd)
    v._store = st
    return v


def _variable_on_cpu(name, shape, initializer):
    """distribute_tools.py:24-38. (Placement is decided by the owner-shard map, not by this helper.)"""
    return _get(name, shape, initializer)


def _add_loss(name, tensor):
    _store().losses.append((name, tensor))


def _variable_with_weight_decay(name, shape, stddev, wd, kind="generic", store_shape=None):
    """distribute_tools.py:41-66: truncated-normal init; adds wd * l2_loss(var) to the 'losses' collection."""
    var = _get(name, shape, TruncatedNormalInitializer(stddev=stddev), kind=kind, store_shape=store_shape, wd=wd)
    if wd is not None and wd != 0.0:
        if var.shadow is not None:
            _add_loss(var.name + "/weight_loss", Fn.l2_loss(var, wd))
        else:
            _add_loss(var.name + "/weight_loss", 0.5 * wd * (var.master ** 2).sum())
    return var


def _pad8(c):
    return (c + 7) // 8 * 8


def _same_pad_nchw(x, k, s):
    H, W = x.shape[2], x.shape[3]
    from ..ops.raw import same_pad
    OH, pt = same_pad(H, k[0], s[0])
    OW, pl = same_pad(W, k[1], s[1])
    pb = max((OH - 1) * s[0] + k[0] - H, 0) - pt
    pr = max((OW - 1) * s[1] + k[1] - W, 0) - pl
    return (pl, pr, pt, pb)


def _stride_hw(stride):
    if isinstance(stride, int):
        return stride, stride
    if len(stride) == 4:
        return int(stride[1]), int(stride[2])
    return int(stride[0]), int(stride[1])


def _prep_cuda_input(x, cin_store):
    """Accept fp32 / uint8 / bf16 NHWC; return bf16 with channels padded to the filter's stored Cin."""
    if x.dtype != torch.bfloat16 or x.shape[-1] != cin_store:
        if x.dtype == torch.bfloat16:
            x = F.pad(x, (0, cin_store - x.shape[-1]))
        else:
            # uint8 images are mapped to [0, 1] (tf.image.convert_image_dtype semantics); floats pass unchanged
            x = Fn.convert_input(x, cin_store, scale=(1.0 / 255.0) if x.dtype == torch.uint8 else 1.0)
    return x.contiguous()


def _chan_pad(x):
    """CUDA activations for the memory-bound kernels: bf16, contiguous, channel count a multiple of 8 (16-byte
    vectors). Layers whose channel count is not get zero channels appended here and cut off again after the op
    (a slow path for unusual widths; zeros are neutral for max / average pooling windows taken per channel, LRN sums,
    batch statistics of their own column, ReLU and add). Returns (tensor, logical channel count)."""
    C = x.shape[-1]
    if x.dtype != torch.bfloat16:
        x = x.to(torch.bfloat16)
    if C % 8:
        x = F.pad(x, (0, _pad8(C) - C))
    return x.contiguous(), C


def _chan_cut(y, C):
    return y if y.shape[-1] == C else y[..., :C]


def _conv_impl(x, w, b, stride, relu, fork=False, emit_stats=False):
    sh, sw = _stride_hw(stride)
    KH, KW = w.shape[0], w.shape[1]
    if x.is_cuda:
        assert sh == sw, "anisotropic strides are not supported on the CUDA path"
        x = _prep_cuda_input(x, w.store_shape[2])
        if fork or emit_stats:
            want_stats = emit_stats and _training  # inference uses the running statistics: nothing to emit
            if fork or want_stats:
                y, x_fwd, part = Fn.conv2d(x, w, b, sh, relu, fork=fork, emit_stats=want_stats)
            else:
                y, x_fwd, part = Fn.conv2d(x, w, b, sh, relu), None, None
            if w.store_shape[3] != w.shape[3]:
                y, part = y[..., :w.shape[3]], None
            return y, x_fwd, part
        y = Fn.conv2d(x, w, b, sh, relu)
        # an output-channel count that is not a multiple of 8 is stored padded (zero filters); slice it off
        return y if w.store_shape[3] == w.shape[3] else y[..., :w.shape[3]]
    if fork or emit_stats:
        return _conv_impl(x, w, b, stride, relu), (x if fork else None), None
    xn = x.float().permute(0, 3, 1, 2)
    if x.dtype == torch.uint8:
        xn = xn / 255.0
    xn = F.pad(xn, _same_pad_nchw(xn, (KH, KW), (sh, sw)))
    y = F.conv2d(xn, w.master.permute(3, 2, 0, 1), b.master if b is not None else None, stride=(sh, sw))
    if relu:
        y = F.relu(y)
    return _q(y.permute(0, 2, 3, 1))


def conv(layer_name, x, in_channels, out_channels, kernel_size=[3, 3], stride=[1, 1, 1, 1]):
    """conv2d(SAME) + bias + ReLU; W ~ trunc-normal(0.05), wd=0.0 -- distribute_tools.py:69-79."""
    with variable_scope(layer_name):
        shape = [kernel_size[0], kernel_size[1], in_channels, out_channels]
        store = [kernel_size[0], kernel_size[1], _pad8(in_channels), _pad8(out_channels)] if x.is_cuda else None
        w = _variable_with_weight_decay("weights", shape, stddev=5e-2, wd=0.0, kind="conv_w", store_shape=store)
        b = _get("biases", [out_channels], ConstantInitializer(0.0), kind="bias",
                 store_shape=[_pad8(out_channels)] if x.is_cuda else None)
        return _conv_impl(x, w, b, stride, relu=True)


def conv_eval(layer_name, x, in_channels, out_channels, kernel_size=[3, 3], stride=[1, 1, 1, 1]):
    """Same op with AUTO_REUSE (distribute_tools.py:81-91): variables are shared by name with `conv`."""
    return conv(layer_name, x, in_channels, out_channels, kernel_size, stride)


def conv_nonacti(layer_name, x, in_channels, out_channels, kernel_size=[3, 3], stride=[1, 1, 1, 1], use_bias=True,
                 fork_input=False, emit_stats=False):
    """conv2d + bias, Xavier init, variables `weights_nonacti` / `biases_nonacti` -- distribute_tools.py:93-106.
    `use_bias=False` (extension) drops the bias for convolutions that feed a batch norm, where it is redundant.
    With `fork_input` or `emit_stats` (extensions) the result is (y, x', part). x' (fork_input): feed it to the other
    consumer of x (a residual shortcut) and the two gradients of x are summed inside this layer's data-gradient
    kernel instead of by a separate pass. part (emit_stats): per-tile batch-norm statistics of y computed in the
    convolution's epilogue, for batch_norm_channel(..., stats_part=part); None where the kernel cannot."""
    with variable_scope(layer_name):
        shape = [kernel_size[0], kernel_size[1], in_channels, out_channels]
        store = [kernel_size[0], kernel_size[1], _pad8(in_channels), _pad8(out_channels)] if x.is_cuda else None
        w = _get("weights_nonacti", shape, XavierInitializer(), kind="conv_w", store_shape=store)
        b = _get("biases_nonacti", [out_channels], ConstantInitializer(0.0), kind="bias",
                 store_shape=[_pad8(out_channels)] if x.is_cuda else None) if use_bias else None
        return _conv_impl(x, w, b, stride, relu=False, fork=fork_input, emit_stats=emit_stats)


def conv_nonacti_eval(layer_name, x, in_channels, out_channels, kernel_size=[3, 3], stride=[1, 1, 1, 1]):
    """distribute_tools.py:108-121."""
    return conv_nonacti(layer_name, x, in_channels, out_channels, kernel_size, stride)


def acti_layer(x):
    """ReLU -- distribute_tools.py:123-128."""
    if x.is_cuda:
        xp, C = _chan_pad(x)
        return _chan_cut(Fn.relu(xp), C)
    return _q(F.relu(x))


def deconv(layer_name, x, in_channels, out_channels, output_shape=[32, 224, 224, 64], kernel_size=[3, 3],
           stride=[1, 1, 1, 1]):
    """conv2d_transpose(SAME), no bias / activation, Xavier init -- distribute_tools.py:131-143.

    As in TF the filter is [kh, kw, in_channels, out_channels] where dim 2 (`in_channels` here) is the number of
    OUTPUT channels of the transposed conv and dim 3 must equal x's channel count."""
    sh, sw = _stride_hw(stride)
    with variable_scope(layer_name):
        # CUDA: both channel counts are stored padded to multiples of 8 (zero filters), like the conv layers
        store = [kernel_size[0], kernel_size[1], _pad8(in_channels), _pad8(out_channels)] if x.is_cuda else None
        w = _get("weights", [kernel_size[0], kernel_size[1], in_channels, out_channels], XavierInitializer(),
                 kind="deconv_w", store_shape=store)
        out_hw = (int(output_shape[1]), int(output_shape[2]))
        if x.is_cuda:
            assert sh == sw
            y = Fn.conv2d_transpose(_prep_cuda_input(x, w.store_shape[3]), w, out_hw, sh)
            return y if w.store_shape[2] == in_channels else y[..., :in_channels]
        # oracle: adjoint of the SAME conv (out_hw -> x spatial). conv_transpose2d without padding yields the
        # gradient w.r.t. the zero-padded input; cropping by TF's asymmetric (top/left) padding gives the result.
        from ..ops.raw import same_pad
        xn = x.float().permute(0, 3, 1, 2)
        _, pt = same_pad(out_hw[0], kernel_size[0], sh)
        _, pl = same_pad(out_hw[1], kernel_size[1], sw)
        full = F.conv_transpose2d(xn, w.master.permute(3, 2, 0, 1), stride=(sh, sw))
        need_h, need_w = pt + out_hw[0], pl + out_hw[1]
        full = F.pad(full, (0, max(need_w - full.shape[3], 0), 0, max(need_h - full.shape[2], 0)))
        return _q(full[:, :, pt:need_h, pl:need_w].permute(0, 2, 3, 1))


def deconv_eval(layer_name, x, in_channels, out_channels, output_shape=[32, 224, 224, 64], kernel_size=[3, 3],
                stride=[1, 1, 1, 1]):
    """distribute_tools.py:145-157."""
    return deconv(layer_name, x, in_channels, out_channels, output_shape, kernel_size, stride)


def pool(layer_name, x, kernel=[1, 2, 2, 1], stride=[1, 2, 2, 1], is_max_pool=True):
    """max / avg pooling, SAME -- distribute_tools.py:160-165."""
    k = (int(kernel[1]), int(kernel[2])) if len(kernel) == 4 else tuple(kernel)
    s = _stride_hw(stride)
    if x.is_cuda:
        xp, C = _chan_pad(x)
        return _chan_cut(Fn.pool2d(xp, k, s, is_max_pool), C)
    xn = x.float().permute(0, 3, 1, 2)
    pad = _same_pad_nchw(xn, k, s)
    if is_max_pool:
        y = F.max_pool2d(F.pad(xn, pad, value=float("-inf")), k, s)
    else:
        num = F.avg_pool2d(F.pad(xn, pad), k, s, divisor_override=1)
        cnt = F.avg_pool2d(F.pad(torch.ones_like(xn), pad), k, s, divisor_override=1)
        y = num / cnt
    return _q(y.permute(0, 2, 3, 1))


def batch_norm(x):
    """The reference's literal batch norm: moments over axis 0 only, no scale / offset, eps 1e-3, no moving
    averages -- distribute_tools.py:168-180."""
    epsilon = 1e-3
    if x.is_cuda:
        xp, C = _chan_pad(x)
        return _chan_cut(Fn.batch_norm_train(xp, rows=x.shape[0], eps=epsilon), C)
    xf = x.float()
    mean = xf.mean(0, keepdim=True)
    var = xf.var(0, unbiased=False, keepdim=True)
    return _q((xf - mean) * torch.rsqrt(var + epsilon))


def _bn_vars(layer_name, C, cuda=False):
    # on CUDA a channel count that is not a multiple of 8 is stored padded (the kernels work on 8-channel vectors);
    # checkpoints keep the logical shape
    ss = [_pad8(C)] if cuda and C % 8 else None
    with variable_scope(layer_name):
        beta = _get("beta", [C], ConstantInitializer(0.0), kind="bias", store_shape=ss)
        gamma = _get("gamma", [C], ConstantInitializer(1.0), kind="bias", store_shape=ss)
        rm = _get("moving_mean", [C], ConstantInitializer(0.0), trainable=False, store_shape=ss)
        rv = _get("moving_variance", [C], ConstantInitializer(1.0), trainable=False, store_shape=ss)
    return gamma, beta, rm, rv


def batch_norm_channel(layer_name, x, relu=False, residual=None, eps=1e-3, momentum=0.1, stats_part=None,
                       residual_bn=None, residual_part=None, _logical_c=None):
    """Standard per-channel batch norm with affine parameters and running statistics (needed by the ResNet
    configs of BASELINE.json; optionally fused with a residual add and ReLU). `stats_part`: per-tile statistics of x
    from the convolution that produced it (conv_nonacti(..., emit_stats=True)); saves the statistics pass.
    `residual_bn` (a layer name): the residual is the RAW output of a projection shortcut's convolution and goes
    through its own batch norm (variables under that name) inside the same pass -- y = act(bn(x) + bn_r(residual));
    `residual_part`: that convolution's epilogue statistics."""
    C_log = _logical_c or x.shape[-1]
    if x.is_cuda and x.shape[-1] % 8:
        # unusual width: zero channels appended for the 8-channel vector kernels, cut off again below
        x = _chan_pad(x)[0]
        residual = _chan_pad(residual)[0] if residual is not None else None
        stats_part = residual_part = None
    C = x.shape[-1]
    gamma, beta, rm, rv = _bn_vars(layer_name, C_log, x.is_cuda)
    rows = x.numel() // C
    if residual_bn is not None:
        assert residual is not None
        gamma_r, beta_r, rm_r, rv_r = _bn_vars(residual_bn, C_log, x.is_cuda)
        if x.is_cuda and _training:
            y = Fn.batch_norm_dual_train(x.contiguous(), residual.contiguous(), rows, gamma, beta, rm.master,
                                         rv.master, gamma_r, beta_r, rm_r.master, rv_r.master, eps, momentum, relu,
                                         stats_part, residual_part)
            return _chan_cut(y, C_log)
        # eval / CPU: normalise the shortcut on its own, then the plain fused path
        residual = batch_norm_channel(residual_bn, residual, relu=False, eps=eps, momentum=momentum,
                                      stats_part=residual_part, _logical_c=C_log)
        if x.is_cuda and residual.shape[-1] != C:
            residual = _chan_pad(residual)[0]
    if x.is_cuda:
        x = x.contiguous()
        if _training:
            y = Fn.batch_norm_train(x, rows, gamma, beta, rm.master, rv.master, eps, momentum, residual, relu,
                                    part=stats_part)
        else:
            y = Fn.batch_norm_eval(x, rows, gamma, beta, rm.master, rv.master, eps, residual, relu)
        return _chan_cut(y, C_log)
    xf = x.float().reshape(rows, C)
    y = F.batch_norm(xf, rm.master.detach(), rv.master.detach(), gamma.master, beta.master, _training, momentum, eps)
    y = y.reshape(x.shape)
    if residual is not None:
        y = y + residual.float()
    return _q(F.relu(y) if relu else y)


def batch_norm_relu_pool(bn_name, pool_name, x, kernel=[1, 3, 3, 1], stride=[1, 2, 2, 1], eps=1e-3, momentum=0.1,
                         stats_part=None):
    """max_pool(relu(batch_norm_channel(x))) -- the ResNet stem. On CUDA in training mode one kernel pools over the
    normalised values computed on the fly (3x3 / 2x2 windows); elsewhere it is the composition of the three layers.
    Variables are those of batch_norm_channel(bn_name)."""
    C = x.shape[-1]
    k = (int(kernel[1]), int(kernel[2])) if len(kernel) == 4 else tuple(kernel)
    s = _stride_hw(stride)
    if x.is_cuda and _training and C % 8 == 0 and k in ((3, 3), (2, 2)):
        gamma, beta, rm, rv = _bn_vars(bn_name, C, True)
        return Fn.batch_norm_relu_pool_train(x, x.numel() // C, gamma, beta, rm.master, rv.master, k, s, eps, momentum,
                                             stats_part)
    y = batch_norm_channel(bn_name, x, relu=True, eps=eps, momentum=momentum, stats_part=stats_part)
    return pool(pool_name, y, kernel=kernel, stride=stride)


def norm(name, x, lsize=4):
    """tf.nn.lrn(depth_radius=lsize, bias=1, alpha=0.001/9, beta=0.75) -- distribute_tools.py:183-187. The reference
    discards the result and returns x unchanged; that behaviour is kept behind --lrn_compat_noop."""
    if _flags.FLAGS.lrn_compat_noop:
        return x
    alpha, beta, bias = 0.001 / 9.0, 0.75, 1.0
    if x.is_cuda:
        xp, C = _chan_pad(x)
        return _chan_cut(Fn.lrn(xp, lsize, bias, alpha, beta), C)
    n = 2 * lsize + 1
    return _q(F.local_response_norm(x.float().permute(0, 3, 1, 2), n, alpha=alpha * n, beta=beta,
                                    k=bias).permute(0, 2, 3, 1))


def _fc_impl(layer_name, x, out_nodes, relu, out_fp32=False):
    size = 1
    for d in x.shape[1:]:
        size *= int(d)
    with variable_scope(layer_name):
        w = _get("weights", [size, out_nodes], XavierInitializer(), kind="fc_w")
        b = _get("biases", [out_nodes], ConstantInitializer(0.0), kind="bias")
        if x.is_cuda:
            flat = x.reshape(-1, size)
            if flat.dtype != torch.bfloat16:
                flat = flat.to(torch.bfloat16)
            return Fn.linear(flat.contiguous(), w, b, relu=relu, out_fp32=out_fp32)
        y = x.float().reshape(-1, size) @ w.master + b.master
        # fp32 logits on CUDA (out_fp32): only the gradient is rounded there
        return _q(F.relu(y) if relu else y, fwd=not out_fp32)


def FC_layer(layer_name, x, out_nodes):
    """flatten + matmul + bias + ReLU, Xavier init -- distribute_tools.py:190-208."""
    return _fc_impl(layer_name, x, out_nodes, relu=True)


def fc_nonacti(layer_name, x, out_nodes):
    """Linear head without activation (logits layer of the CIFAR-10 tutorial / ResNet); fp32 logits on CUDA."""
    return _fc_impl(layer_name, x, out_nodes, relu=False, out_fp32=True)


def global_avg_pool(x):
    """[N,H,W,C] -> [N,C] mean over the spatial dims (ResNet head)."""
    H, W = x.shape[1], x.shape[2]
    if x.is_cuda:
        xp, C = _chan_pad(x)
        return _chan_cut(Fn.pool2d(xp, (H, W), (H, W), False).reshape(x.shape[0], xp.shape[3]), C)
    return _q(x.float().mean(dim=(1, 2)))


def add_relu(a, b, relu=True):
    if a.is_cuda:
        ap, C = _chan_pad(a)
        return _chan_cut(Fn.add(ap, _chan_pad(b)[0], relu), C)
    y = a.float() + b.float()
    return _q(F.relu(y) if relu else y)


def weight(kernel_shape, is_uniform=True):
    """distribute_tools.py:211-216."""
    return _get("weights", list(kernel_shape), XavierInitializer())


def bias(bias_shape):
    """distribute_tools.py:219-224."""
    return _get("biases", list(bias_shape), ConstantInitializer(0.0), kind="bias")
