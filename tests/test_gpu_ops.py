"""Numerics of the memory-bound sm_100a kernels vs plain PyTorch fp32 references of the same op."""
import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def raw():
    from distributed_tensorflow_framework_b200.ops import raw as r
    return r


def rnd(*shape):
    return torch.randn(*shape, device="cuda").to(torch.bfloat16)


def close(got, ref, tol=2e-2):
    got, ref = got.float(), ref.float()
    err = (got - ref).abs().max().item()
    scale = ref.abs().max().item() + 1e-6
    assert not torch.isnan(got).any()
    assert err / scale < tol, (err, scale)


def test_convert_input(raw):
    x = torch.rand(4, 32, 32, 3, device="cuda")
    y = raw.convert_input(x, scale=2.0, shift=-1.0)
    assert y.shape == (4, 32, 32, 8)
    close(y[..., :3], x * 2 - 1)
    assert (y[..., 3:] == 0).all()
    u = torch.randint(0, 255, (2, 8, 8, 3), device="cuda", dtype=torch.uint8)
    close(raw.convert_input(u, scale=1 / 255.)[..., :3], u.float() / 255.)


def test_relu_add(raw):
    x = rnd(4, 16, 16, 64)
    y = raw.relu(x)
    close(y, torch.relu(x.float()))
    dy = rnd(4, 16, 16, 64)
    close(raw.relu_bwd(dy, y), dy.float() * (y.float() > 0))
    close(raw.add(x, dy, relu=True), torch.relu(x.float() + dy.float()))


@pytest.mark.parametrize("R,N", [(1000, 64), (4096, 384), (333, 192), (128, 4096), (77, 8)])
def test_bias_grad(raw, R, N):
    dy = rnd(R, N)
    y = rnd(R, N)
    close(raw.bias_grad(dy), dy.float().sum(0), 1e-2)
    dz, db = raw.relu_bwd_bias_grad(dy, y)
    ref = dy.float() * (y.float() > 0)
    close(dz, ref)
    close(db, ref.sum(0), 1e-2)


@pytest.mark.parametrize("k,s,is_max", [((3, 3), (2, 2), True), ((2, 2), (2, 2), True), ((3, 3), (2, 2), False),
                                         ((2, 2), (2, 2), False), ((3, 3), (1, 1), True)])
def test_pool(raw, k, s, is_max):
    n, H, W, C = 3, 17, 16, 64
    x = rnd(n, H, W, C)
    OH, OW, pt, pl = raw.pool_geometry(H, W, k, s)
    pb = max((OH - 1) * s[0] + k[0] - H, 0) - pt
    pr = max((OW - 1) * s[1] + k[1] - W, 0) - pl
    xf = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    if is_max:
        ref = F.max_pool2d(F.pad(xf, (pl, pr, pt, pb), value=float("-inf")), k, s)
    else:
        num = F.avg_pool2d(F.pad(xf, (pl, pr, pt, pb)), k, s, divisor_override=1)
        cnt = F.avg_pool2d(F.pad(torch.ones_like(xf), (pl, pr, pt, pb)), k, s, divisor_override=1)
        ref = num / cnt
    y, idx = raw.pool_fwd(x, k, s, is_max)
    close(y, ref.permute(0, 2, 3, 1))
    dy = rnd(n, OH, OW, C)
    ref.backward(dy.float().permute(0, 3, 1, 2))
    close(raw.pool_bwd(idx, dy, x.shape, k, s, is_max), xf.grad.permute(0, 2, 3, 1))


def test_lrn(raw):
    x = rnd(2, 8, 8, 64) * 3
    r, bias, alpha, beta = 4, 1.0, 0.001 / 9.0 * 500, 0.75  # exaggerated alpha so that the window matters
    xf = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    n = 2 * r + 1
    ref = F.local_response_norm(xf, n, alpha=alpha * n, beta=beta, k=bias)  # torch divides alpha by n, TF does not
    close(raw.lrn_fwd(x, r, bias, alpha, beta), ref.permute(0, 2, 3, 1))
    dy = rnd(2, 8, 8, 64)
    ref.backward(dy.float().permute(0, 3, 1, 2))
    close(raw.lrn_bwd(x, dy, r, bias, alpha, beta), xf.grad.permute(0, 2, 3, 1), 3e-2)


@pytest.mark.parametrize("R,N,relu,with_res", [(2048, 64, True, True), (1024, 128, True, False), (512, 256, False, False),
                                               (64, 8 * 8 * 16, False, False)])
def test_batch_norm(raw, R, N, relu, with_res):
    x = rnd(R, N) * 2 + 0.5
    res = rnd(R, N) if with_res else None
    gamma = torch.rand(N, device="cuda") + 0.5
    beta = torch.randn(N, device="cuda")
    rm, rv = torch.zeros(N, device="cuda"), torch.ones(N, device="cuda")
    xf = x.float().requires_grad_(True)
    g2, b2 = gamma.clone().requires_grad_(True), beta.clone().requires_grad_(True)
    rm_ref, rv_ref = rm.clone(), rv.clone()
    ref = F.batch_norm(xf, rm_ref, rv_ref, g2, b2, training=True, momentum=0.1, eps=1e-3)
    resf = res.float().requires_grad_(True) if with_res else None
    if with_res:
        ref = ref + resf
    if relu:
        ref = torch.relu(ref)
    y, mean, rstd, scale, shift = raw.bn_train_fwd(x, gamma, beta, rm, rv, 1e-3, 0.1, res, relu)
    close(y, ref)
    close(rm, rm_ref, 1e-2)
    close(rv, rv_ref, 1e-2)
    dy = rnd(R, N)
    ref.backward(dy.float())
    dx, dz, dgamma, dbeta = raw.bn_bwd(dy, y, x, mean, rstd, gamma, relu)
    close(dx, xf.grad, 3e-2)
    close(dgamma, g2.grad, 2e-2)
    close(dbeta, b2.grad, 2e-2)
    if with_res:
        close(dz, resf.grad)
    if relu and with_res:  # the forward's ReLU bitmask (1 bit per element) replaces y in the backward
        out = raw.bn_train_fwd(x, gamma, beta, rm.clone(), rv.clone(), 1e-3, 0.1, res, relu, want_mask=True)
        assert torch.equal(out[0], y) and out[5].dtype == torch.uint8 and tuple(out[5].shape) == (R, N // 8)
        bits = ((out[5].long().unsqueeze(-1) >> torch.arange(8, device="cuda")) & 1).reshape(R, N).bool()
        assert torch.equal(bits, y > 0)
        dx3, dz3, dg3, db3 = raw.bn_bwd(dy, out[5], x, mean, rstd, gamma, relu)
        assert torch.equal(dz3, dz) and torch.equal(dx3, dx)
        close(dg3, g2.grad, 2e-2)
    if relu and not with_res:  # mask recomputed from x: same result without reading y / writing dz
        dx2, dz2, dg2, db2 = raw.bn_bwd(dy, None, x, mean, rstd, gamma, relu, mask_scale=scale, mask_shift=shift)
        assert dz2 is None
        close(dx2, xf.grad, 3e-2)
        close(dg2, g2.grad, 2e-2)
    close(raw.bn_eval_fwd(x, gamma, beta, rm, rv, 1e-3),
          F.batch_norm(x.float(), rm, rv, gamma, beta, False, 0.0, 1e-3))


@pytest.mark.parametrize("B,C,dtype", [(128, 10, torch.float32), (128, 10, torch.bfloat16), (37, 1000, torch.float32)])
def test_softmax_xent(raw, B, C, dtype):
    logits = (torch.randn(B, C, device="cuda") * 3).to(dtype)
    labels = torch.randint(0, C, (B,), device="cuda")
    lf = logits.float().requires_grad_(True)
    ref = F.cross_entropy(lf, labels)
    ref.backward()
    for _ in range(2):  # twice: the scratch counter must be left clean
        loss, dl = raw.softmax_xent(logits, labels)
        close(loss, ref.detach().view(1), 1e-3)
        close(dl, lf.grad, 2e-2 if dtype == torch.bfloat16 else 1e-4)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_mse_l2(raw, dtype):
    p = torch.randn(4, 32, 32, 8, device="cuda").to(dtype)
    t = torch.randn(4, 32, 32, 8, device="cuda").to(dtype)
    pf = p.float().requires_grad_(True)
    ref = F.mse_loss(pf, t.float())
    ref.backward()
    for _ in range(2):
        loss, dp = raw.mse(p, t)
        close(loss, ref.detach().view(1), 1e-3)
        close(dp, pf.grad, 2e-2 if dtype == torch.bfloat16 else 1e-4)
    w = torch.randn(100003, device="cuda")
    close(raw.l2_loss(w, 0.5), (0.25 * (w * w).sum()).view(1), 1e-4)


def test_top1(raw):
    logits = torch.randn(100, 10, device="cuda")
    labels = torch.randint(0, 10, (100,), device="cuda")
    assert raw.top1_hits(logits, labels).item() == (logits.argmax(1) == labels).sum().item()


@pytest.mark.parametrize("kind", ["sgd", "momentum", "adam"])
def test_optimizers(raw, kind):
    n = 10007
    lr = 0.01 if kind == "adam" else 0.1
    b1, b2, eps = 0.9, 0.999, 1e-8
    p = torch.randn(n, device="cuda")
    pr = p.clone()
    mr, vr = torch.zeros_like(p), torch.zeros_like(p)
    m = torch.zeros(n, device="cuda")
    v = torch.zeros(n, device="cuda")
    shadow = torch.empty(n, device="cuda", dtype=torch.bfloat16)
    for step in range(1, 4):
        g = torch.randn(n, device="cuda")
        ge = g * 0.5
        # plain PyTorch fp32 statement of the TF update rules (tf.train.{GradientDescent,Momentum,Adam}Optimizer)
        if kind == "sgd":
            pr = pr - lr * ge
        elif kind == "momentum":
            mr = b1 * mr + ge
            pr = pr - lr * mr
        else:
            mr = b1 * mr + (1 - b1) * ge
            vr = b2 * vr + (1 - b2) * ge * ge
            lr_t = lr * (1 - b2 ** step) ** 0.5 / (1 - b1 ** step)
            pr = pr - lr_t * mr / (vr.sqrt() + eps)
        raw.optim_step(kind, p, g, m, v, shadow, lr, b1=b1, b2=b2, eps=eps, gscale=0.5, step=step)
    close(p, pr, 1e-5)
    close(shadow, p)


def test_init(raw):
    t = torch.empty(1 << 20, device="cuda")
    raw.init_(t, "truncated_normal", 0.0, 0.05, seed=7)
    assert t.abs().max().item() <= 0.1 + 1e-6
    assert abs(t.mean().item()) < 1e-3
    assert abs(t.std().item() - 0.05 * 0.8796) < 2e-3  # std of a normal truncated at 2 sigma
    raw.init_(t, "uniform", -0.5, 0.5, seed=8)
    assert t.min().item() >= -0.5 and t.max().item() < 0.5 and abs(t.mean().item()) < 2e-3
    t2 = torch.empty_like(t)
    raw.init_(t2, "uniform", -0.5, 0.5, seed=8)
    assert torch.equal(t, t2)  # counter-based: reproducible
    raw.init_(t, "constant", 0.25)
    assert (t == 0.25).all()


def test_pool_over_batch_norm_relu_matches_the_composition(raw):
    """dtf_pool_fwd_affine: max_pool(relu(x * scale + shift)) computed inside the pooling kernel equals pooling the
    separately normalised tensor (the ResNet stem path), values and arg-max routing."""
    x = rnd(4, 30, 30, 64)
    scale = torch.randn(64, device="cuda")
    shift = torch.randn(64, device="cuda") * 0.3
    ref_in = torch.relu(x.float() * scale + shift).to(torch.bfloat16)
    y_ref, idx_ref = raw.pool_fwd(ref_in, (3, 3), (2, 2), True)
    y, idx = raw.pool_fwd(x, (3, 3), (2, 2), True, pre_scale=scale, pre_shift=shift)
    assert torch.equal(y, y_ref)
    dy = rnd(*y.shape)
    d1 = raw.pool_bwd(idx, dy, tuple(x.shape), (3, 3), (2, 2), True)
    d2 = raw.pool_bwd(idx_ref, dy, tuple(x.shape), (3, 3), (2, 2), True)
    # ties (several window elements equal after bf16 rounding / all clipped to 0) may pick another element
    assert (d1 != d2).float().mean() < 0.15
    assert torch.allclose(d1.float().sum((1, 2)), d2.float().sum((1, 2)), rtol=2e-2, atol=2e-1)


def test_softmax_xent_out_of_range_label_is_nan_not_a_wild_read(raw):
    logits = torch.randn(16, 10, device="cuda")
    labels = torch.randint(0, 10, (16,), device="cuda")
    labels[3] = 10 ** 12  # far outside the row: must not be used as an index
    labels[5] = -7
    loss, dl = raw.softmax_xent(logits, labels)
    torch.cuda.synchronize()
    assert bool(torch.isnan(loss).all())
    sm = torch.softmax(logits, 1) / 16
    assert torch.allclose(dl[3], sm[3], atol=1e-6) and torch.allclose(dl[5], sm[5], atol=1e-6)
    good = torch.ones(16, dtype=torch.bool, device="cuda")
    good[3] = good[5] = False
    ref = torch.softmax(logits, 1)
    ref[torch.arange(16), labels.clamp(0, 9)] -= 1
    assert torch.allclose(dl[good], (ref / 16)[good], atol=1e-6)
    loss2, _ = raw.softmax_xent(logits, labels.clamp(0, 9))  # the scratch counter was left clean
    assert bool(torch.isfinite(loss2).all())
