"""The allocation-free batch-norm forward, called through the C ABI.

Some hosts of this library cannot allocate device memory themselves -- they only hold buffers lent to
them by another runtime, or they run where cudaMalloc's device-wide stall is not acceptable -- so
``dcgan_bn_forward_static`` keeps its per-channel sums in an array built into the library.  This test
calls it twice in a row (different inputs, so stale sums from the first call would show) and compares
each result with the ordinary ``dcgan_bn_stats`` + ``dcgan_bn_apply`` pair fed a caller-owned buffer.
It also checks that an over-size request is refused.  It lives in the last-collected file so that a
``-x`` run has already exercised every other test before reaching it, and it sets up everything it
needs itself, so it also passes when run alone in a fresh process.
"""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu

_ACT_LRELU = 2


def _entry(lib):
    fn = lib.dcgan_bn_forward_static
    fn.argtypes = [ctypes.c_void_p] * 7 + [ctypes.c_int] * 3 + [ctypes.c_float] * 2 + [ctypes.c_int, ctypes.c_void_p]
    fn.restype = ctypes.c_int
    return fn


def test_bn_forward_static_matches_caller_buffer_path(cuda_lib):
    from distributed_tensorflow_for_dcgan_b200.ops import kernels as K
    fn = _entry(cuda_lib)
    groups, rows, C = 2, 2048, 128
    gamma = 1 + 0.1 * torch.randn(C, device="cuda")
    beta = 0.1 * torch.randn(C, device="cuda")
    stream = torch.cuda.current_stream().cuda_stream
    for call in range(2):
        x = ((torch.randn(groups * rows, C, device="cuda") * (1.0 + call)) + 0.3 * call).to(torch.bfloat16)
        # ordinary path: the caller owns the sums buffer
        sums = torch.zeros(groups, 2, C, device="cuda")
        saved_ref = torch.zeros(groups, 2, C, device="cuda")
        em_ref = torch.zeros(groups, C, device="cuda")
        ev_ref = torch.zeros(groups, C, device="cuda")
        y_ref = torch.empty_like(x)
        K.bn_stats(x, sums, groups)
        K.bn_apply(x, y_ref, sums, gamma, beta, saved_ref, em_ref, ev_ref, groups, "lrelu")
        # allocation-free path
        saved = torch.zeros(groups, 2, C, device="cuda")
        em = torch.zeros(groups, C, device="cuda")
        ev = torch.zeros(groups, C, device="cuda")
        y = torch.empty_like(x)
        rc = fn(x.data_ptr(), y.data_ptr(), gamma.data_ptr(), beta.data_ptr(), saved.data_ptr(), em.data_ptr(),
                ev.data_ptr(), groups, rows, C, 1e-5, 0.9, _ACT_LRELU, stream)
        assert rc == 0
        torch.cuda.synchronize()
        den = y_ref.float().abs().max().item()
        assert (y.float() - y_ref.float()).abs().max().item() <= 1e-2 * den, f"call {call}"
        assert torch.allclose(saved, saved_ref, rtol=2e-3, atol=1e-5), f"call {call}"
        assert torch.allclose(em, em_ref, rtol=2e-3, atol=1e-6) and torch.allclose(ev, ev_ref, rtol=2e-3, atol=1e-6)
    K.check_device_error()


def test_bn_forward_static_rejects_oversize(cuda_lib):
    fn = _entry(cuda_lib)
    groups, rows, C = 8, 8, 2048  # 8*2*2048 = 32768 floats > 16384
    x = torch.zeros(groups * rows, C, dtype=torch.bfloat16, device="cuda")
    y = torch.empty_like(x)
    g = torch.ones(C, device="cuda")
    b = torch.zeros(C, device="cuda")
    rc = fn(x.data_ptr(), y.data_ptr(), g.data_ptr(), b.data_ptr(), None, None, None, groups, rows, C, 1e-5, 0.9,
            0, torch.cuda.current_stream().cuda_stream)
    assert rc == -4  # DCGAN_ERR_LIMIT
