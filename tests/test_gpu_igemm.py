"""Numerics of the tcgen05 GEMM / implicit-GEMM convolution kernels vs PyTorch fp32 references."""
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
    assert not torch.isnan(got).any()
    err = (got - ref).abs().max().item()
    assert err / (ref.abs().max().item() + 1e-6) < tol


@pytest.mark.parametrize("M,N,K", [(128, 128, 64), (200, 192, 136), (128, 384, 4096), (1024, 64, 576), (128, 10, 192),
                                   (4096, 256, 128)])
def test_gemm(raw, M, N, K):
    a, b, bias = rnd(M, K), rnd(K, N), torch.randn(N, device="cuda")
    ref = a.float() @ b.float()
    close(raw.gemm(a, b), ref)
    close(raw.gemm(a, b, bias, relu=True, out_fp32=True), torch.relu(ref + bias), 1e-4)
    at = a.t().contiguous()
    close(raw.gemm(at, b, a_transposed=True), ref)
    acc = torch.zeros(M, N, device="cuda")
    raw.gemm(at, b, a_transposed=True, out=acc, accumulate=True)
    close(acc, ref, 1e-4)


CONVS = [(2, 16, 16, 64, 64, 3, 3, 1), (2, 56, 56, 64, 64, 3, 3, 1), (3, 28, 28, 128, 128, 3, 3, 1),
         (2, 14, 14, 256, 128, 3, 3, 1), (2, 17, 19, 64, 192, 3, 3, 1), (2, 12, 12, 64, 64, 5, 5, 1), (1, 40, 150, 64, 64, 3, 3, 1), (2, 64, 64, 128, 64, 3, 3, 1), (2, 7, 7, 64, 64, 3, 3, 1), (4, 32, 32, 8, 64, 5, 5, 1), (2, 28, 28, 128, 256, 3, 3, 2),
         (2, 14, 14, 256, 512, 1, 1, 1), (2, 56, 56, 64, 128, 1, 1, 2), (2, 30, 30, 16, 32, 7, 7, 2),
         (3, 17, 19, 32, 64, 3, 3, 2)]


@pytest.mark.parametrize("n,H,W,Ci,Co,KH,KW,s", CONVS)
def test_conv_family(raw, n, H, W, Ci, Co, KH, KW, s):
    x = rnd(n, H, W, Ci)
    w = (torch.randn(KH, KW, Ci, Co, device="cuda") * 0.05).to(torch.bfloat16)
    bias = torch.randn(Co, device="cuda")
    xf = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    wf = w.float().permute(3, 2, 0, 1).requires_grad_(True)
    OH, pt = raw.same_pad(H, KH, s)
    OW, pl = raw.same_pad(W, KW, s)
    pb = max((OH - 1) * s + KH - H, 0) - pt
    pr = max((OW - 1) * s + KW - W, 0) - pl
    yref = F.conv2d(F.pad(xf, (pl, pr, pt, pb)), wf, stride=s)
    y = raw.conv_fprop(x, w.reshape(KH * KW * Ci, Co), bias, KH, KW, s, relu=True)
    close(y, torch.relu(yref + bias.view(1, -1, 1, 1)).permute(0, 2, 3, 1))
    dy = rnd(n, OH, OW, Co)
    yref.backward(dy.float().permute(0, 3, 1, 2))
    w_t = w.permute(0, 1, 3, 2).contiguous().reshape(KH * KW * Co, Ci)
    close(raw.conv_dgrad(dy, w_t, (H, W), KH, KW, s), xf.grad.permute(0, 2, 3, 1))
    close(raw.conv_wgrad(x, dy, KH, KW, s).view(KH, KW, Ci, Co), wf.grad.permute(2, 3, 1, 0), 3e-2)


def test_halo_kernels_match_the_gather_kernels(raw):
    """The halo-tile fprop / dgrad / wgrad kernels against the im2col-gather kernels on the same problem."""
    n, H, W, Ci, Co, k = 3, 28, 28, 128, 64, 3
    x, dy = rnd(n, H, W, Ci), rnd(n, H, W, Co)
    w = (torch.randn(k, k, Ci, Co, device="cuda") * 0.05).to(torch.bfloat16)
    w2 = w.reshape(k * k * Ci, Co)
    wt = w.permute(0, 1, 3, 2).contiguous().reshape(k * k * Co, Ci)
    assert raw._lib.lib().dtf_conv_halo_supported(H, W, Ci, k, k, Co) == 1
    assert raw._lib.lib().dtf_conv_wgrad_halo_supported(H, W, Ci, k, k, Co) == 1
    got = (raw.conv_fprop(x, w2, None, k, k, 1), raw.conv_dgrad(dy, wt, (H, W), k, k, 1), raw.conv_wgrad(x, dy, k, k, 1))
    raw.USE_HALO = False
    try:
        ref = (raw.conv_fprop(x, w2, None, k, k, 1), raw.conv_dgrad(dy, wt, (H, W), k, k, 1),
               raw.conv_wgrad(x, dy, k, k, 1))
    finally:
        raw.USE_HALO = True
    for g, r in zip(got, ref):
        close(g, r, 1e-2)


def test_persistent_gemm_matches_one_tile_per_cta(raw):
    a, b = rnd(1000, 320), rnd(320, 512)
    bias = torch.randn(512, device="cuda")
    y_p = raw.gemm(a, b, bias, relu=True)
    prev = raw._lib.lib().dtf_set_persistent_gemm(0)
    try:
        y_1 = raw.gemm(a, b, bias, relu=True)
    finally:
        raw._lib.lib().dtf_set_persistent_gemm(prev)
    assert torch.equal(y_p, y_1)


def test_strided_pixel_select_and_scatter(raw):
    x = rnd(2, 9, 10, 16)
    y = raw.subsample(x, 2)
    assert torch.equal(y, x[:, ::2, ::2, :])
    back = raw.upsample_zero(y, (9, 10), 2)
    ref = torch.zeros_like(x)
    ref[:, ::2, ::2, :] = y
    assert torch.equal(back, ref)


PACKED = [(2, 224, 224, 8, 3, 64, 7, 2), (3, 24, 24, 8, 3, 64, 5, 1), (2, 40, 36, 8, 3, 64, 3, 1), (2, 56, 56, 128, 128, 128, 3, 2),
          (2, 14, 14, 512, 512, 512, 3, 2), (2, 33, 31, 64, 64, 128, 3, 2), (2, 16, 16, 64, 64, 8, 3, 1)]


@pytest.mark.parametrize("n,H,W,Cx,Cu,Co,K,s", PACKED)
def test_packed_window_convs(raw, n, H, W, Cx, Cu, Co, K, s):
    """Strided / few-channel convolutions as pack + stride-1 halo convolution (fprop, dgrad, wgrad) vs PyTorch."""
    x = rnd(n, H, W, Cx)
    x[..., Cu:] = 0
    w = (torch.randn(K, K, Cx, Co, device="cuda") * 0.05).to(torch.bfloat16)
    xf = x.float().permute(0, 3, 1, 2).requires_grad_(True)
    wf = w.float().permute(3, 2, 0, 1).requires_grad_(True)
    OH, pt = raw.same_pad(H, K, s)
    OW, pl = raw.same_pad(W, K, s)
    pb = max((OH - 1) * s + K - H, 0) - pt
    pr = max((OW - 1) * s + K - W, 0) - pl
    yref = F.conv2d(F.pad(xf, (pl, pr, pt, pb)), wf, stride=s)
    assert raw._pack_plan(H, W, Cx, Cu, K, K, s) is not None or Co % 64 != 0
    y = raw.conv_fprop(x, w.reshape(K * K * Cx, Co), None, K, K, s, c_used=Cu)
    close(y, yref.permute(0, 2, 3, 1))
    dy = rnd(n, OH, OW, Co)
    yref.backward(dy.float().permute(0, 3, 1, 2))
    w_t = w.permute(0, 1, 3, 2).contiguous().reshape(K * K * Co, Cx)
    close(raw.conv_dgrad(dy, w_t, (H, W), K, K, s), xf.grad.permute(0, 2, 3, 1))
    dw = raw.conv_wgrad(x, dy, K, K, s, c_used=Cu).view(K, K, Cx, Co)
    close(dw[:, :, :Cu], wf.grad.permute(2, 3, 1, 0)[:, :, :Cu], 3e-2)
    raw.USE_PACK = False
    try:
        y0 = raw.conv_fprop(x, w.reshape(K * K * Cx, Co), None, K, K, s)
    finally:
        raw.USE_PACK = True
    close(y, y0, 1e-2)


@pytest.mark.parametrize("M,N,K", [(4096, 256, 64), (1000, 72, 136), (128, 512, 1024)])
def test_gemm_addend(raw, M, N, K):
    """act(A @ B + bias + addend): the addend is read in the (staged) epilogue of the tcgen05 GEMM."""
    a, b, add, bias = rnd(M, K), rnd(K, N), rnd(M, N), torch.randn(N, device="cuda")
    ref = a.float() @ b.float() + add.float()
    close(raw.gemm(a, b, addend=add), ref)
    close(raw.gemm(a, b, bias, relu=True, addend=add), torch.relu(ref + bias))


def test_conv_fork_sums_both_gradients_of_x():
    """conv2d(fork=True) forwards x to a second consumer; its backward adds that consumer's gradient inside the
    data-gradient kernel (1x1: GEMM epilogue; 3x3: add kernel). Compared with plain autograd accumulation."""
    from distributed_tensorflow_framework_b200.engine import functional as Fn, variables as V
    from distributed_tensorflow_framework_b200.engine.variables import XavierInitializer
    for k in (1, 3):
        store = V.VariableStore(device="cuda")
        w = store.get_variable("w%d" % k, [k, k, 64, 128], XavierInitializer(), kind="conv_w")
        x = rnd(2, 16, 16, 64).requires_grad_(True)
        g1, g2 = rnd(2, 16, 16, 128), rnd(2, 16, 16, 64)
        y, xf, _ = Fn.conv2d(x, w, None, 1, False, fork=True)
        (y * g1).sum().backward(retain_graph=True, inputs=[x])
        gx_main = x.grad.clone()
        x.grad = None
        ((y * g1).sum() + (xf * g2).sum()).backward(inputs=[x])
        close(x.grad, gx_main.float() + g2.float())


@pytest.mark.parametrize("n,H,C,Co,k,s", [(4, 28, 128, 128, 3, 1), (5, 7, 64, 72, 3, 1), (3, 56, 64, 64, 3, 1),
                                          (2, 30, 64, 128, 3, 2), (2, 64, 8, 64, 7, 2)])
def test_conv_epilogue_statistics(raw, n, H, C, Co, k, s):
    x = rnd(n, H, H, C)
    w = (torch.randn(k, k, C, Co, device="cuda") * 0.05).to(torch.bfloat16).reshape(k * k * C, Co)
    st = {}
    y = raw.conv_fprop(x, w, None, k, k, s, stats=st, c_used=min(C, 3) if C == 8 else None)
    assert "part" in st
    tot = st["part"].sum(0)
    yf = y.float().reshape(-1, Co)
    assert torch.allclose(tot[0], yf.sum(0), rtol=2e-2, atol=0.5)
    assert torch.allclose(tot[1], (yf * yf).sum(0), rtol=2e-2, atol=0.5)


@pytest.mark.parametrize("M,N,K", [(802816 // 8, 256, 64), (50176, 1024, 256), (12544, 1024, 512), (3136, 64, 256),
                                   (1000, 128, 64), (200704 // 4, 512, 128)])
def test_gemm_epilogue_statistics(raw, M, N, K):
    """Batch-norm statistics from the dense GEMM epilogue (1x1 convolutions): the per-warp partial rows add up to the
    column sums / sums of squares of the stored bf16 output, the output itself is unchanged, and the 1x1 convolution
    entry point (plain and strided) hands them to batch norm."""
    a = rnd(M, K)
    b = (torch.randn(K, N, device="cuda") * 0.1).to(torch.bfloat16)
    st = {}
    y = raw.gemm(a, b, stats=st)
    assert "part" in st and st["part"].shape[1:] == (2, N) and 1 <= st["part"].shape[0] <= 296
    assert torch.equal(y, raw.gemm(a, b))
    tot = st["part"].double().sum(0)
    yf = y.double()
    assert torch.allclose(tot[0], yf.sum(0), rtol=1e-4, atol=1e-2)
    assert torch.allclose(tot[1], (yf * yf).sum(0), rtol=1e-4, atol=1e-2)
    st2 = {}
    raw.gemm(a, b, relu=True, stats=st2)  # a fused ReLU changes what is stored: no statistics offered
    assert "part" not in st2


def test_conv1x1_statistics_feed_batch_norm(raw):
    x = rnd(8, 28, 28, 128)
    w = (torch.randn(1, 1, 128, 256, device="cuda") * 0.05).to(torch.bfloat16).reshape(128, 256)
    for stride in (1, 2):
        st = {}
        y = raw.conv_fprop(x, w, None, 1, 1, stride, stats=st)
        assert "part" in st
        y2 = y.reshape(-1, 256)
        gamma, beta = torch.rand(256, device="cuda") + 0.5, torch.randn(256, device="cuda")
        o1 = raw.bn_train_fwd(y2, gamma, beta, None, None, 1e-3, 0.1, part=st["part"])
        o2 = raw.bn_train_fwd(y2, gamma, beta, None, None, 1e-3, 0.1)
        close(o1[1], o2[1], 1e-3)  # mean
        close(o1[2], o2[2], 1e-3)  # rstd
        close(o1[0], o2[0], 1e-2)


def test_gather_conv_with_more_than_65535_row_tiles(raw):
    """8.5 M output pixels = 66 k row tiles: tiles are linearised in grid.x (grid.y would overflow)."""
    n, H, W, C, Co = 170, 224, 224, 8, 8
    x = rnd(n, H, W, C)
    w = (torch.randn(3, 3, C, Co, device="cuda") * 0.1).to(torch.bfloat16)
    raw.USE_PACK = False
    try:
        y = raw.conv_fprop(x, w.reshape(9 * C, Co), None, 3, 3, 1)
    finally:
        raw.USE_PACK = True
    for img in (0, n - 1):  # the last image lives in the tiles beyond 65535
        ref = F.conv2d(x[img:img + 1].float().permute(0, 3, 1, 2), w.float().permute(3, 2, 0, 1), padding=1)
        close(y[img:img + 1], ref.permute(0, 2, 3, 1))


def test_tap_masks_do_not_change_the_result(raw):
    """Skipping the structurally zero weight tiles of a space-to-depth packed stride-2 convolution is exact."""
    n, H, C, Co = 3, 28, 128, 256
    x, dy = rnd(n, H, H, C), rnd(n, H // 2, H // 2, Co)
    w = (torch.randn(3, 3, C, Co, device="cuda") * 0.05).to(torch.bfloat16)
    wt = w.permute(0, 1, 3, 2).contiguous().reshape(9 * Co, C)
    got = (raw.conv_fprop(x, w.reshape(9 * C, Co), None, 3, 3, 2), raw.conv_dgrad(dy, wt, (H, H), 3, 3, 2))
    raw.USE_TAP_MASKS = False
    try:
        ref = (raw.conv_fprop(x, w.reshape(9 * C, Co), None, 3, 3, 2), raw.conv_dgrad(dy, wt, (H, H), 3, 3, 2))
    finally:
        raw.USE_TAP_MASKS = True
    for g, r in zip(got, ref):
        assert torch.equal(g, r)


BOUNDARY_CONVS = [(2, 7, 7, 64, 64, 3, 3, 1), (1, 5, 9, 64, 128, 3, 3, 1), (2, 17, 19, 64, 192, 3, 3, 1),
                  (3, 17, 19, 32, 64, 3, 3, 2), (2, 13, 13, 128, 128, 3, 3, 2), (2, 9, 9, 64, 64, 5, 5, 1),
                  (2, 15, 15, 256, 64, 1, 1, 2), (1, 33, 33, 8, 64, 7, 7, 2)]


@pytest.mark.parametrize("n,H,W,Ci,Co,KH,KW,s", BOUNDARY_CONVS)
def test_conv_family_against_cpu_fp32_oracle(raw, n, H, W, Ci, Co, KH, KW, s):
    """Boundary-heavy shapes (small / odd maps where most outputs touch the TF-SAME padding, asymmetric for stride 2)
    against torch's CPU fp32 convolution -- not cuDNN on the same GPU -- with a per-element mixed bound
    |err| <= atol + rtol |ref| (bf16 inputs, fp32 accumulation, bf16 output rounding)."""
    from distributed_tensorflow_framework_b200.ops.raw import same_pad
    g = torch.Generator().manual_seed(H * 1000 + W * 10 + KH)
    x = torch.randn(n, H, W, Ci, generator=g).to(torch.bfloat16)
    w = (torch.randn(KH, KW, Ci, Co, generator=g) * 0.1).to(torch.bfloat16)
    OH, pt = same_pad(H, KH, s)
    OW, pl = same_pad(W, KW, s)
    pb = max((OH - 1) * s + KH - H, 0) - pt
    pr = max((OW - 1) * s + KW - W, 0) - pl
    xc = F.pad(x.float().permute(0, 3, 1, 2), (pl, pr, pt, pb)).requires_grad_(True)
    wc = w.float().permute(3, 2, 0, 1).clone().requires_grad_(True)
    ref = F.conv2d(xc, wc, stride=s)
    dy = torch.randn(n, OH, OW, Co, generator=g).to(torch.bfloat16)
    ref.backward(dy.float().permute(0, 3, 1, 2))
    ref_y = ref.detach().permute(0, 2, 3, 1)
    ref_dx = xc.grad[:, :, pt:pt + H, pl:pl + W].permute(0, 2, 3, 1)
    ref_dw = wc.grad.permute(2, 3, 1, 0).reshape(KH * KW * Ci, Co)

    def bound(got, want, atol, rtol):
        got, want = got.float().cpu(), want.float()
        bad = (got - want).abs() > atol + rtol * want.abs()
        assert not bool(bad.any()), (int(bad.sum()), float((got - want).abs().max()), float(want.abs().max()))

    xg, wg, dyg = x.cuda(), w.cuda(), dy.cuda()
    w2 = wg.reshape(KH * KW * Ci, Co)
    wt = wg.permute(0, 1, 3, 2).contiguous().reshape(KH * KW * Co, Ci)
    scale = float(ref_y.abs().max())
    bound(raw.conv_fprop(xg, w2, None, KH, KW, s), ref_y, 1e-2 * scale, 1e-2)
    bound(raw.conv_dgrad(dyg, wt, (H, W), KH, KW, s), ref_dx, 1e-2 * float(ref_dx.abs().max()), 1e-2)
    dw = torch.zeros(KH * KW * Ci, Co, device="cuda")
    raw.conv_wgrad(xg, dyg, KH, KW, s, out=dw)
    bound(dw, ref_dw, 2e-3 * float(ref_dw.abs().max()), 2e-3)
