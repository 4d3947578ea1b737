"""torch.autograd.Function wrappers around the sm_100a kernels (CUDA path only; the CPU path uses plain torch ops).

Parameters enter a Function through their bf16 shadow leaf (requires_grad=True) so that autograd schedules the
backward even for the first layer; the weight gradient is NOT returned to autograd: the wgrad / bias-grad kernels
accumulate straight into the fp32 gradient arena (`Variable.grad`), then the store is told that the gradient is
ready so the engine can start the bucket's collective while backward continues (replaces
optimizer.compute_gradients of /root/reference/distribute_tower.py:26-27).
"""
import torch

from ..ops import raw

bf16 = torch.bfloat16


def _ready(var):
    st = getattr(var, "_store", None)
    if st is not None:
        st.notify_grad_ready(var)


class UseVariableFn(torch.autograd.Function):
    """Variable.value() on CUDA: forward hands out the bf16 shadow; backward accumulates into the fp32 gradient arena
    (the variable may be used several times in one forward pass) and announces the gradient once the last
    outstanding use of this step has delivered its part."""

    @staticmethod
    def forward(ctx, shadow, var):
        ctx.var = var
        st = getattr(var, "_store", None)
        uses = getattr(st, "_bridge_uses", None) if st is not None else None
        if uses is not None:
            uses[var.name] = uses.get(var.name, 0) + 1
        return shadow.view_as(shadow)

    @staticmethod
    def backward(ctx, g):
        var = ctx.var
        var.grad.add_(g.reshape(var.grad.shape))
        st = getattr(var, "_store", None)
        uses = getattr(st, "_bridge_uses", None) if st is not None else None
        left = 0
        if uses is not None and var.name in uses:
            uses[var.name] -= 1
            left = uses[var.name]
        if left <= 0:
            _ready(var)
        return None, None


def use_variable(var):
    return UseVariableFn.apply(var.shadow, var)


class ConvertInputFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, c_pad, scale, shift):
        return raw.convert_input(x, c_pad, scale, shift)

    @staticmethod
    def backward(ctx, dy):
        return None, None, None, None


def convert_input(x, c_pad=None, scale=1.0, shift=0.0):
    return ConvertInputFn.apply(x, c_pad, scale, shift)


class ConvFn(torch.autograd.Function):
    """conv2d(SAME) + bias (+ ReLU) on the implicit-GEMM tcgen05 kernel."""

    @staticmethod
    def forward(ctx, x, w_shadow, b_shadow, wvar, bvar, stride, relu, fork=False, emit_stats=False):
        KH, KW, Ci, Co = wvar.store_shape
        w2d = w_shadow.detach().view(KH * KW * Ci, Co)
        bias = bvar.master if bvar is not None else None
        keep = {}
        stats = {} if emit_stats else None
        y = raw.conv_fprop(x, w2d, bias, KH, KW, stride, relu=relu, c_used=wvar.shape[2], keep=keep, stats=stats)
        ctx.wvar, ctx.bvar, ctx.stride, ctx.relu = wvar, bvar, stride, relu
        # strided / few-channel layers ran on a packed copy of x: keep that for the filter gradient instead of x
        ctx.packed = "packed" in keep
        ctx.x_shape = tuple(x.shape)
        ctx.save_for_backward(keep["packed"] if ctx.packed else x, y if relu else None)
        ctx.x_needs_grad = x.requires_grad
        ctx.fork = fork
        if fork or emit_stats:
            # second output: x itself, for the other consumer of x (the shortcut branch). Its gradient comes back
            # into THIS backward and is summed inside the data-gradient kernel, so autograd never runs a separate
            # accumulation pass over the block input.
            # third output: per-tile batch-norm statistics of y computed in the kernel's epilogue (or None).
            part = stats.get("part") if stats is not None else None
            if part is not None:
                ctx.mark_non_differentiable(part)
            return y, (x.view_as(x) if fork else None), part
        return y

    @staticmethod
    def backward(ctx, dy, d_alias=None, _d_part=None):
        x, y = ctx.saved_tensors
        wvar, bvar = ctx.wvar, ctx.bvar
        KH, KW, Ci, Co = wvar.store_shape
        if dy is None:  # only the forwarded copy of x was used
            return d_alias, None, None, None, None, None, None, None, None
        dy = dy.contiguous()
        n, OH, OW, _ = dy.shape
        dy2 = dy.view(-1, Co)
        if ctx.relu:
            if bvar is not None:
                dz2, _ = raw.relu_bwd_bias_grad(dy2, y.view(-1, Co), bvar.grad)
            else:
                dz2 = raw.relu_bwd(dy2, y.view(-1, Co))
        else:
            dz2 = dy2
            if bvar is not None:
                raw.bias_grad(dy2, bvar.grad)
        dz = dz2.view(n, OH, OW, Co)
        if ctx.packed:
            raw.conv_wgrad(None, dz, KH, KW, ctx.stride, out=wvar.grad.view(KH * KW * Ci, Co), c_used=wvar.shape[2],
                           packed=x, x_shape=ctx.x_shape)
        else:
            raw.conv_wgrad(x, dz, KH, KW, ctx.stride, out=wvar.grad.view(KH * KW * Ci, Co), c_used=wvar.shape[2])
        _ready(wvar)
        if bvar is not None:
            _ready(bvar)
        dx = None
        if ctx.x_needs_grad:
            dx = raw.conv_dgrad(dz, wvar.shadow_t.view(KH * KW * Co, Ci), ctx.x_shape[1:3], KH, KW, ctx.stride,
                                addend=d_alias)
        return dx, None, None, None, None, None, None, None, None


def conv2d(x, wvar, bvar, stride=1, relu=False, fork=False, emit_stats=False):
    """With `fork` or `emit_stats` the result is (y, x', part): x' (fork) is x for a second consumer whose gradient is
    summed inside this op's backward; part (emit_stats) holds per-tile batch-norm statistics of y, or None."""
    return ConvFn.apply(x, wvar.shadow, bvar.shadow if bvar is not None else None, wvar, bvar, stride, relu, fork,
                        emit_stats)


class DeconvFn(torch.autograd.Function):
    """conv2d_transpose(SAME): forward is the data-gradient GEMM; filter [KH,KW,Cout,Cin] (TF layout)."""

    @staticmethod
    def forward(ctx, x, w_shadow, wvar, out_hw, stride):
        KH, KW, Cout, Cin = wvar.store_shape
        y = raw.conv_dgrad(x, wvar.shadow_t.view(KH * KW * Cin, Cout), out_hw, KH, KW, stride)
        ctx.wvar, ctx.stride = wvar, stride
        ctx.save_for_backward(x)
        ctx.x_needs_grad = x.requires_grad
        return y

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        wvar = ctx.wvar
        KH, KW, Cout, Cin = wvar.store_shape
        dy = dy.contiguous()
        # dW[kh,kw,co,ci] = sum_p dy[p*s - pad + k, co] * x[p, ci]: the wgrad kernel with (dy, x) in the roles (x, dy)
        raw.conv_wgrad(dy, x, KH, KW, ctx.stride, out=wvar.grad.view(KH * KW * Cout, Cin))
        # dx reads the plain shadow, which the bucket's collective rewrites once the gradient is announced: compute it
        # BEFORE _ready (conv / linear read `shadow_t`, which is only refreshed in finish_step)
        dx = None
        if ctx.x_needs_grad:
            dx = raw.conv_fprop(dy, wvar.shadow.detach().view(KH * KW * Cout, Cin), None, KH, KW, ctx.stride)
        _ready(wvar)
        return dx, None, None, None, None


def conv2d_transpose(x, wvar, out_hw, stride=1):
    return DeconvFn.apply(x, wvar.shadow, wvar, tuple(out_hw), stride)


class LinearFn(torch.autograd.Function):
    """act(x @ W + b), W stored [in, out] (TF FC_layer layout)."""

    @staticmethod
    def forward(ctx, x, w_shadow, b_shadow, wvar, bvar, relu, out_fp32):
        w = w_shadow.detach()
        ag = getattr(wvar, "ag", None)
        if ag is not None:  # parameter-server pull fused into this GEMM: weight tiles come from their owners
            y = raw.gemm_ag(x, ag, False, bvar.master if bvar is not None else None, relu=relu, out_fp32=out_fp32)
        else:
            y = raw.gemm(x, w, bvar.master if bvar is not None else None, relu=relu, out_fp32=out_fp32)
        ctx.wvar, ctx.bvar, ctx.relu = wvar, bvar, relu
        ctx.save_for_backward(x, y if relu else None)
        ctx.x_needs_grad = x.requires_grad
        return y

    @staticmethod
    def backward(ctx, dy):
        x, y = ctx.saved_tensors
        wvar, bvar = ctx.wvar, ctx.bvar
        dy = dy.contiguous()
        if dy.dtype != bf16:
            dy = dy.to(bf16)
        N = dy.shape[1]
        if N % 8 == 0:
            if ctx.relu:
                if bvar is not None:
                    dz, _ = raw.relu_bwd_bias_grad(dy, y, bvar.grad)
                else:
                    dz = raw.relu_bwd(dy, y)
            else:
                dz = dy
                if bvar is not None:
                    raw.bias_grad(dy, bvar.grad)
        else:  # tiny, unaligned heads (e.g. 10 logits): plain torch elementwise is fine here
            dz = dy * (y > 0) if ctx.relu else dy
            if bvar is not None:
                bvar.grad.add_(dz.float().sum(0))
        raw.gemm(x, dz, a_transposed=True, out=wvar.grad, accumulate=True)  # dW = x^T dz
        _ready(wvar)
        if bvar is not None:
            _ready(bvar)
        dx = None
        if ctx.x_needs_grad:
            ag = getattr(wvar, "ag", None)
            if ag is not None:
                dx = raw.gemm_ag(dz, ag, True)
            else:
                dx = raw.gemm(dz, wvar.shadow_t)  # dz @ W^T, W^T stored [out, in]
        return dx, None, None, None, None, None, None


def linear(x, wvar, bvar, relu=False, out_fp32=False):
    return LinearFn.apply(x, wvar.shadow, bvar.shadow if bvar is not None else None, wvar, bvar, relu, out_fp32)


class PoolFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, k, s, is_max):
        y, idx = raw.pool_fwd(x, k, s, is_max, want_idx=x.requires_grad)
        ctx.k, ctx.s, ctx.is_max, ctx.in_shape = k, s, is_max, tuple(x.shape)
        ctx.save_for_backward(idx)
        return y

    @staticmethod
    def backward(ctx, dy):
        (idx,) = ctx.saved_tensors
        return raw.pool_bwd(idx, dy.contiguous(), ctx.in_shape, ctx.k, ctx.s, ctx.is_max), None, None, None


def pool2d(x, k=(2, 2), s=(2, 2), is_max=True):
    return PoolFn.apply(x, tuple(k), tuple(s), is_max)


class LRNFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x, radius, bias, alpha, beta):
        ctx.args = (radius, bias, alpha, beta)
        ctx.save_for_backward(x)
        return raw.lrn_fwd(x, radius, bias, alpha, beta)

    @staticmethod
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        return raw.lrn_bwd(x, dy.contiguous(), *ctx.args), None, None, None, None


def lrn(x, radius=4, bias=1.0, alpha=0.001 / 9.0, beta=0.75):
    return LRNFn.apply(x, radius, bias, alpha, beta)


class ReLUFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, x):
        y = raw.relu(x)
        ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        (y,) = ctx.saved_tensors
        return raw.relu_bwd(dy.contiguous(), y)


def relu(x):
    return ReLUFn.apply(x)


class AddFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, a, b, relu):
        y = raw.add(a, b, relu)
        ctx.relu = relu
        if relu:
            ctx.save_for_backward(y)
        return y

    @staticmethod
    def backward(ctx, dy):
        dy = dy.contiguous()
        if ctx.relu:
            (y,) = ctx.saved_tensors
            dy = raw.relu_bwd(dy, y)
        return dy, dy, None


def add(a, b, relu=False):
    return AddFn.apply(a, b, relu)


class BatchNormFn(torch.autograd.Function):
    """Training-mode batch norm over the rows of x viewed as [R, N] (+ residual, + ReLU)."""

    @staticmethod
    def forward(ctx, x, res, g_shadow, gvar, bvar, run_mean, run_var, rows, eps, momentum, relu, part=None):
        shape = x.shape
        x2 = x.reshape(rows, -1)
        res2 = res.reshape(rows, -1) if res is not None else None
        gamma = gvar.master if gvar is not None else None
        beta = bvar.master if bvar is not None else None
        if part is not None and tuple(part.shape[1:]) != (2, x2.shape[1]):
            part = None
        # without a residual the ReLU mask is a function of x alone: keep the folded coefficients instead of y;
        # with one, the forward writes the decisions as a bitmask (1 bit per element) and the backward reads that
        ctx.mask_from_x = relu and res is None
        want_mask = relu and res is not None and x2.shape[1] % 8 == 0
        out = raw.bn_train_fwd(x2, gamma, beta, run_mean, run_var, eps, momentum, res2, relu, part=part,
                               want_mask=want_mask)
        y2, mean, rstd, scale, shift = out[:5]
        ctx.gvar, ctx.bvar, ctx.relu, ctx.rows, ctx.has_res = gvar, bvar, relu, rows, res is not None
        if ctx.mask_from_x:
            ctx.save_for_backward(x2, None, mean, rstd, scale, shift)
        elif want_mask:
            ctx.save_for_backward(x2, out[5], mean, rstd, None, None)
        else:
            ctx.save_for_backward(x2, y2 if relu else None, mean, rstd, None, None)
        return y2.view(shape)

    @staticmethod
    def backward(ctx, dy):
        x2, y2, mean, rstd, msc, msh = ctx.saved_tensors
        gvar, bvar = ctx.gvar, ctx.bvar
        shape = dy.shape
        dy2 = dy.contiguous().view(ctx.rows, -1)
        # the coefficient kernel adds d gamma / d beta straight into the gradient arena
        dx2, dz2, _, _ = raw.bn_bwd(dy2, y2, x2, mean, rstd, gvar.master if gvar is not None else None, ctx.relu,
                                    dgamma_acc=gvar.grad if gvar is not None else None,
                                    dbeta_acc=bvar.grad if bvar is not None else None, mask_scale=msc, mask_shift=msh)
        if gvar is not None:
            _ready(gvar)
        if bvar is not None:
            _ready(bvar)
        dres = dz2.view(shape) if ctx.has_res else None
        return dx2.view(shape), dres, None, None, None, None, None, None, None, None, None, None


def batch_norm_train(x, rows, gvar=None, bvar=None, run_mean=None, run_var=None, eps=1e-3, momentum=0.1, res=None,
                     relu=False, part=None):
    return BatchNormFn.apply(x, res, gvar.shadow if gvar is not None else None, gvar, bvar, run_mean, run_var, rows,
                             eps, momentum, relu, part)


class BnReluPoolFn(torch.autograd.Function):
    """max_pool(relu(batch_norm(x))): the normalised tensor is produced inside the pooling kernel's window loads and
    never stored (ResNet stem: saves a write and a read of the largest activation of the network)."""

    @staticmethod
    def forward(ctx, x, g_shadow, gvar, bvar, rm, rv, rows, eps, momentum, k, s, part):
        x = x.contiguous()
        x2 = x.reshape(rows, -1)
        mean, rstd, scale, shift = raw.bn_stats_coeffs(x2, gvar.master, bvar.master, rm, rv, eps, momentum, part)
        y, idx = raw.pool_fwd(x, k, s, True, want_idx=True, pre_scale=scale, pre_shift=shift)
        ctx.gvar, ctx.bvar, ctx.rows, ctx.k, ctx.s, ctx.in_shape = gvar, bvar, rows, k, s, tuple(x.shape)
        ctx.save_for_backward(x2, idx, mean, rstd, scale, shift)
        return y

    @staticmethod
    def backward(ctx, dy):
        x2, idx, mean, rstd, scale, shift = ctx.saved_tensors
        gvar, bvar = ctx.gvar, ctx.bvar
        d_bn = raw.pool_bwd(idx, dy.contiguous(), ctx.in_shape, ctx.k, ctx.s, True)
        dx2, _, _, _ = raw.bn_bwd(d_bn.view(ctx.rows, -1), None, x2, mean, rstd, gvar.master, True,
                                  dgamma_acc=gvar.grad, dbeta_acc=bvar.grad, mask_scale=scale, mask_shift=shift)
        _ready(gvar)
        _ready(bvar)
        return (dx2.view(ctx.in_shape),) + (None,) * 11


def batch_norm_relu_pool_train(x, rows, gvar, bvar, rm, rv, k, s, eps=1e-3, momentum=0.1, part=None):
    return BnReluPoolFn.apply(x, gvar.shadow, gvar, bvar, rm, rv, rows, eps, momentum, tuple(k), tuple(s), part)


class DualBatchNormFn(torch.autograd.Function):
    """y = relu(bn_a(x) + bn_b(xs)): the closing batch norm of a residual block whose shortcut is a projection
    (convolution + batch norm). One apply pass reads both raw convolution outputs; the shortcut's normalised tensor is
    never written. Backward: dz = dy * mask (bitmask from the forward), then each batch norm's own reduce + apply."""

    @staticmethod
    def forward(ctx, x, xs, ga_shadow, gs_shadow, gvar, bvar, rm, rv, gvar_s, bvar_s, rm_s, rv_s, rows, eps, momentum,
                relu, part, part_s):
        shape = x.shape
        x2, xs2 = x.reshape(rows, -1), xs.reshape(rows, -1)
        mean, rstd, scale, shift = raw.bn_stats_coeffs(x2, gvar.master, bvar.master, rm, rv, eps, momentum, part)
        mean_s, rstd_s, scale_s, shift_s = raw.bn_stats_coeffs(xs2, gvar_s.master, bvar_s.master, rm_s, rv_s, eps,
                                                               momentum, part_s)
        y2, mask = raw.bn_apply_dual(x2, xs2, scale, shift, scale_s, shift_s, relu)
        ctx.vars = (gvar, bvar, gvar_s, bvar_s)
        ctx.relu, ctx.rows = relu, rows
        ctx.save_for_backward(x2, xs2, mask, mean, rstd, mean_s, rstd_s)
        return y2.view(shape)

    @staticmethod
    def backward(ctx, dy):
        x2, xs2, mask, mean, rstd, mean_s, rstd_s = ctx.saved_tensors
        gvar, bvar, gvar_s, bvar_s = ctx.vars
        shape = dy.shape
        dy2 = dy.contiguous().view(ctx.rows, -1)
        dx2, dz2, _, _ = raw.bn_bwd(dy2, mask if ctx.relu else None, x2, mean, rstd, gvar.master, ctx.relu,
                                    dgamma_acc=gvar.grad, dbeta_acc=bvar.grad)
        _ready(gvar)
        _ready(bvar)
        dxs2, _, _, _ = raw.bn_bwd(dz2, None, xs2, mean_s, rstd_s, gvar_s.master, False, dgamma_acc=gvar_s.grad,
                                   dbeta_acc=bvar_s.grad)
        _ready(gvar_s)
        _ready(bvar_s)
        return (dx2.view(shape), dxs2.view(shape)) + (None,) * 16


def batch_norm_dual_train(x, xs, rows, gvar, bvar, rm, rv, gvar_s, bvar_s, rm_s, rv_s, eps=1e-3, momentum=0.1,
                          relu=True, part=None, part_s=None):
    return DualBatchNormFn.apply(x, xs, gvar.shadow, gvar_s.shadow, gvar, bvar, rm, rv, gvar_s, bvar_s, rm_s, rv_s, rows,
                                 eps, momentum, relu, part, part_s)


def batch_norm_eval(x, rows, gvar, bvar, run_mean, run_var, eps=1e-3, res=None, relu=False):
    shape = x.shape
    y = raw.bn_eval_fwd(x.reshape(rows, -1), gvar.master if gvar is not None else None,
                        bvar.master if bvar is not None else None, run_mean, run_var, eps,
                        res.reshape(rows, -1) if res is not None else None, relu)
    return y.view(shape)


class SoftmaxXentFn(torch.autograd.Function):
    """Fused mean sparse softmax cross-entropy: forward also produces dlogits (scaled in backward)."""

    @staticmethod
    def forward(ctx, logits, labels):
        loss, dl = raw.softmax_xent(logits, labels, 1.0, want_grad=True)
        ctx.save_for_backward(dl)
        return loss.view(())

    @staticmethod
    def backward(ctx, g):
        (dl,) = ctx.saved_tensors
        return dl * g.to(dl.dtype), None


def softmax_xent(logits, labels):
    return SoftmaxXentFn.apply(logits, labels)


class MSEFn(torch.autograd.Function):
    @staticmethod
    def forward(ctx, pred, target):
        loss, dp = raw.mse(pred, target, 1.0, want_grad=True)
        ctx.save_for_backward(dp)
        return loss.view(())

    @staticmethod
    def backward(ctx, g):
        (dp,) = ctx.saved_tensors
        return dp * g.to(dp.dtype), None


def mse(pred, target):
    return MSEFn.apply(pred, target)


class L2Fn(torch.autograd.Function):
    """wd * sum(w^2)/2 (tf.nn.l2_loss weight-decay term, distribute_tools.py:63-65).

    The gradient wd * w is added to the gradient arena at FORWARD time (the arena was zeroed by begin_step and the
    filter-gradient kernels accumulate with red.add): an autograd node of this term would run after the layer's own
    backward has already announced the gradient to the bucket collective. The term therefore assumes the total loss is
    the plain sum of the collected losses (coefficient 1), which is what Tower does (distribute_tower.py:133-135).
    In the owner-sharded modes the fp32 master of a filter is current only on its owner, so w is read from the
    replicated bf16 shadow there."""

    @staticmethod
    def forward(ctx, w_shadow, wvar, wd, track):
        st = getattr(wvar, "_store", None)
        replicated = st is None or getattr(st, "masters_replicated", True)
        w = wvar.master if replicated else wvar.shadow.detach().float()
        if track:
            wvar.grad.add_(w, alpha=float(wd))
        return raw.l2_loss(w.contiguous(), wd).view(())

    @staticmethod
    def backward(ctx, g):
        return None, None, None, None


def l2_loss(wvar, wd):
    return L2Fn.apply(wvar.shadow, wvar, wd, torch.is_grad_enabled())
