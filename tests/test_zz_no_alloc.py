"""The allocation-free loss entry point, driven through the C ABI exactly as a foreign caller would drive it.

Who needs it: code that runs where device memory cannot be allocated -- the GPU's memory already belongs to the host
application's own pool, or the call sits in a loop where a cudaMalloc would stall the device, or the host's plug-in
rules forbid allocation after start-up. Such a caller passes only its own input and output buffers; the scratch area
comes from a fixed array inside the library. The test checks the result against the regular entry point (which is
given caller-allocated scratch) on two calls in a row, because the second call meets the first call's leftovers in
the shared array, and checks that an oversized batch is refused. This file sorts last so that a run which stops at
the first failure has already executed every other test.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu


def _ptr(t):
    return ctypes.c_void_p(t.data_ptr()) if t is not None else None


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_noalloc_matches_regular_entry_point_twice(dtype):
    from distributed_tensorflow_framework_b200.ops import _lib
    L = _lib.lib()  # ctypes.CDLL of libdtf_b200_cuda.so with the signatures of include/dtf_b200.h
    assert L.dtf_softmax_xent_noalloc_max_rows() == 65536
    dev = torch.device("cuda", 0)
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    is_bf16 = 1 if dtype == torch.bfloat16 else 0
    tol = 2e-2 if dtype == torch.bfloat16 else 1e-4  # same tolerances as tests/test_gpu_ops.py::test_softmax_xent
    for call, (B, C) in enumerate([(300, 10), (128, 1000)]):  # second call: different size, stale partials present
        torch.manual_seed(call)
        logits = (torch.randn(B, C, device=dev) * 3).to(dtype)
        labels = torch.randint(0, C, (B,), device=dev)
        # regular entry point: the caller allocates the scratch (1 + B floats, counter word zero)
        scratch = torch.zeros(1 + B, device=dev, dtype=torch.float32)
        ref_loss = torch.empty(1, device=dev)
        ref_grad = torch.empty_like(logits)
        rc = L.dtf_softmax_xent(_ptr(logits), _ptr(labels), _ptr(ref_grad), _ptr(ref_loss), _ptr(scratch), B, C, 1.0,
                                is_bf16, stream)
        assert rc == 0
        # allocation-free entry point: inputs and outputs only
        loss = torch.empty(1, device=dev)
        grad = torch.empty_like(logits)
        rc = L.dtf_softmax_xent_noalloc(_ptr(logits), _ptr(labels), _ptr(grad), _ptr(loss), B, C, 1.0, is_bf16, stream)
        assert rc == 0
        torch.cuda.synchronize(dev)
        assert abs(loss.item() - ref_loss.item()) <= 1e-3 * max(1.0, abs(ref_loss.item()))
        err = (grad.float() - ref_grad.float()).abs().max().item()
        assert err <= tol * (ref_grad.float().abs().max().item() + 1e-6)
        # and both agree with plain PyTorch
        want = torch.nn.functional.cross_entropy(logits.float(), labels)
        assert abs(loss.item() - want.item()) <= 1e-3 * max(1.0, abs(want.item()))


def test_noalloc_rejects_oversized_batch():
    from distributed_tensorflow_framework_b200.ops import _lib
    L = _lib.lib()
    dev = torch.device("cuda", 0)
    stream = ctypes.c_void_p(torch.cuda.current_stream(dev).cuda_stream)
    B, C = L.dtf_softmax_xent_noalloc_max_rows() + 1, 4
    logits = torch.zeros(B, C, device=dev)
    labels = torch.zeros(B, device=dev, dtype=torch.int64)
    loss = torch.full((1,), -1.0, device=dev)
    rc = L.dtf_softmax_xent_noalloc(_ptr(logits), _ptr(labels), None, _ptr(loss), B, C, 1.0, 0, stream)
    assert rc == 6  # DTF_ERR_TOO_LARGE
    torch.cuda.synchronize(dev)
    assert loss.item() == -1.0  # nothing was launched
