"""CPU tier: the window-packing plans of ops/raw.py (_PackPlan) restated in plain PyTorch.

A strided / few-channel SAME convolution must equal: pack the input windows into channels, re-index the filter
blocks, run a STRIDE-1 convolution with the plan's taps and pads. The CUDA kernels (pack_window / s2d / block_copy /
halo convolution) implement exactly these index maps; here they are checked without a GPU."""
import pytest
import torch
import torch.nn.functional as F

from distributed_tensorflow_framework_b200.ops import raw


def _pack(x, plan):
    """P[n, j, i, (r*kwn + q)*cs + c] = x[n, rp*j + r, sx*i + q - ox, c] (zero outside), as pack_window_kernel."""
    n = x.shape[0]
    P = torch.zeros(n, plan.Hp, plan.Wq, plan.Cp)
    for r in range(plan.rp):
        for q in range(plan.kwn):
            for j in range(plan.Hp):
                y = plan.rp * j + r
                if y >= plan.H:
                    continue
                for i in range(plan.Wq):
                    xx = plan.sx * i + q - plan.ox
                    if 0 <= xx < plan.W:
                        c0 = (r * plan.kwn + q) * plan.cs
                        P[:, j, i, c0:c0 + plan.C] = x[:, y, xx, :plan.C]
    return P


def _block_copy(src, dst, desc):
    t, nblk, r, c = desc
    for src_off, dst_off, rc in t.tolist():
        rows, cols = rc & 0xffffffff, rc >> 32
        assert (rows, cols) == (r, c)
        s = src.reshape(-1)
        d = dst.reshape(-1)
        for k in range(rows):
            d[dst_off + k * dst.stride(0): dst_off + k * dst.stride(0) + cols] = \
                s[src_off + k * src.stride(0): src_off + k * src.stride(0) + cols]


def _same_ref(x, w, s):
    H, W, K = x.shape[1], x.shape[2], w.shape[0]
    OH, pt = raw.same_pad(H, K, s)
    OW, pl = raw.same_pad(W, K, s)
    pb = max((OH - 1) * s + K - H, 0) - pt
    pr = max((OW - 1) * s + K - W, 0) - pl
    xp = F.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb))
    return F.conv2d(xp, w.permute(3, 2, 0, 1), stride=s).permute(0, 2, 3, 1)


CASES = [  # H, W, Cx, C used, K, stride, expected mode
    (12, 12, 16, 16, 3, 2, "s2d"), (11, 13, 16, 16, 3, 2, "s2d"), (14, 14, 16, 16, 7, 2, "s2d"),
    (16, 16, 8, 3, 7, 2, "full"), (15, 17, 8, 3, 7, 2, "full"), (10, 10, 8, 3, 5, 1, "full"), (9, 9, 8, 3, 3, 1, "full"),
]


@pytest.mark.parametrize("H,W,Cx,C,K,s,mode", CASES)
def test_packed_convolution_equals_the_strided_one(H, W, Cx, C, K, s, mode):
    torch.manual_seed(0)
    Co = 8
    plan = raw._pack_plan(H, W, Cx, C, K, K, s)
    assert plan is not None and plan.mode == mode
    x = torch.randn(2, H, W, Cx)
    x[..., C:] = 0
    w = torch.randn(K, K, Cx, Co)
    ref = _same_ref(x, w, s)
    P = _pack(x, plan)
    Wp = torch.zeros(plan.KH2 * plan.KW2 * plan.Cp, Co)
    _block_copy(w.reshape(K * K * Cx, Co), Wp, plan.descs("w", Co, "cpu"))
    # stride-1 convolution over P with taps (KH2, KW2) and SAME-like pads (pad_h2, pad_w2): output = P's spatial size
    wp4 = Wp.reshape(plan.KH2, plan.KW2, plan.Cp, Co)
    Pp = F.pad(P.permute(0, 3, 1, 2), (plan.pad_w2, plan.KW2 - 1 - plan.pad_w2, plan.pad_h2, plan.KH2 - 1 - plan.pad_h2))
    got = F.conv2d(Pp, wp4.permute(3, 2, 0, 1)).permute(0, 2, 3, 1)
    assert got.shape == ref.shape
    assert torch.allclose(got, ref, atol=1e-4, rtol=1e-4)


def test_filter_gradient_blocks_scatter_back_to_the_filter():
    """'dw' descriptors are the inverse of 'w': packed gradient rows land on the original (tap, channel) rows."""
    plan = raw._pack_plan(12, 12, 16, 16, 3, 3, 2)
    Co = 8
    w = torch.randn(9 * 16, Co)
    Wp = torch.zeros(plan.KH2 * plan.KW2 * plan.Cp, Co)
    _block_copy(w, Wp, plan.descs("w", Co, "cpu"))
    back = torch.zeros_like(w)
    _block_copy(Wp, back, plan.descs("dw", Co, "cpu"))
    assert torch.equal(back, w)


def test_space_to_depth_phase_masks_mark_exactly_the_nonzero_weight_blocks():
    plan = raw._pack_plan(12, 12, 16, 16, 3, 3, 2)
    Co = 4
    Wp = torch.zeros(plan.KH2 * plan.KW2 * plan.Cp, Co)
    _block_copy(torch.ones(9 * 16, Co), Wp, plan.descs("w", Co, "cpu"))
    blocks = Wp.reshape(plan.KH2 * plan.KW2, 4, 16, Co).abs().sum((2, 3)) > 0  # [tap, phase]
    for phase in range(4):
        for tap in range(plan.KH2 * plan.KW2):
            assert bool(blocks[tap, phase]) == bool((plan.phase_mask[phase] >> tap) & 1)
    assert int(blocks.sum()) == 9  # 9 of the 16 (tap, phase) blocks carry weights


def test_plans_are_refused_where_packing_does_not_apply():
    assert raw._pack_plan(14, 14, 64, 64, 1, 1, 2) is None      # 1x1: plain pixel selection + GEMM
    assert raw._pack_plan(14, 14, 64, 64, 3, 3, 1) is None      # stride 1, C % 64 == 0: the halo kernel takes it as is
    assert raw._pack_plan(14, 14, 24, 24, 3, 3, 2) is None      # C not a multiple of 16 and too wide to tap-pack


# ---- the same identity for random geometries (rectangular images and kernels, both modes) ----------------------------
from hypothesis import assume, given, settings, strategies as st  # noqa: E402


def _same_ref_rect(x, w, s):
    H, W, KH, KW = x.shape[1], x.shape[2], w.shape[0], w.shape[1]
    OH, pt = raw.same_pad(H, KH, s)
    OW, pl = raw.same_pad(W, KW, s)
    pb = max((OH - 1) * s + KH - H, 0) - pt
    pr = max((OW - 1) * s + KW - W, 0) - pl
    xp = F.pad(x.permute(0, 3, 1, 2), (pl, pr, pt, pb))
    return F.conv2d(xp, w.permute(3, 2, 0, 1), stride=s).permute(0, 2, 3, 1)


@settings(max_examples=60, deadline=None)
@given(H=st.integers(3, 15), W=st.integers(3, 15), KH=st.integers(1, 7), KW=st.integers(1, 7), s=st.sampled_from([1, 2]),
       geom=st.sampled_from([(16, 16), (32, 32), (8, 1), (8, 3), (8, 4), (16, 5)]), seed=st.integers(0, 1000))
def test_packed_convolution_identity_for_random_geometries(H, W, KH, KW, s, geom, seed):
    Cx, C = geom
    plan = raw._pack_plan(H, W, Cx, C, KH, KW, s)
    assume(plan is not None)
    torch.manual_seed(seed)
    Co = 8
    x = torch.randn(1, H, W, Cx)
    x[..., C:] = 0
    w = torch.randn(KH, KW, Cx, Co)
    ref = _same_ref_rect(x, w, s)
    P = _pack(x, plan)
    Wp = torch.zeros(plan.KH2 * plan.KW2 * plan.Cp, Co)
    _block_copy(w.reshape(KH * KW * Cx, Co), Wp, plan.descs("w", Co, "cpu"))
    wp4 = Wp.reshape(plan.KH2, plan.KW2, plan.Cp, Co)
    Pp = F.pad(P.permute(0, 3, 1, 2), (plan.pad_w2, plan.KW2 - 1 - plan.pad_w2, plan.pad_h2, plan.KH2 - 1 - plan.pad_h2))
    got = F.conv2d(Pp, wp4.permute(3, 2, 0, 1)).permute(0, 2, 3, 1)
    assert got.shape == ref.shape, (plan.mode, got.shape, ref.shape)
    assert torch.allclose(got, ref, atol=1e-4, rtol=1e-4), plan.mode
    # and the filter-gradient descriptors invert the filter descriptors
    back = torch.zeros(KH * KW * Cx, Co)
    _block_copy(Wp, back, plan.descs("dw", Co, "cpu"))
    assert torch.equal(back.reshape(KH, KW, Cx, Co)[:, :, :C], w[:, :, :C])
