"""p2p_instnorm_stats_static: the allocation-free variant of p2p_instnorm_stats (scratch in a file-scope device array of
the library; include/p2p_b200.h).  Called through the C ABI and compared with the existing entry point on the same
inputs, two calls in a row (the second on a different, smaller tensor, so stale scratch contents would show), within
the 1e-5 relative tolerance the fp32 InstanceNorm tests use (tests/test_nets_gpu.py).  Collected last on purpose."""
import ctypes

import pytest
import torch

from cases import relerr
from pix2pixhdaudiosr_b200 import _lib

pytestmark = pytest.mark.gpu
TOL = 1e-5
_DT = {torch.float32: 0, torch.bfloat16: 1}


def _p(t):
    return ctypes.c_void_p(t.data_ptr())


def _stream():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def _both(y, eps=1e-5):
    """(existing entry point, allocation-free entry point) statistics [N,C,2] of an NHWC tensor; outputs are allocated
    here, before the calls, so that nothing is allocated between the two synchronisation points around the new call"""
    lib = _lib.lib()
    N, H, W, C = y.shape
    want = torch.full((N, C, 2), float("nan"), device=y.device)
    got = torch.full((N, C, 2), float("nan"), device=y.device)
    _lib.check(lib.p2p_instnorm_stats(_p(y), _p(want), N, H * W, C, eps, _DT[y.dtype], _stream()))
    torch.cuda.synchronize()
    free0, _ = torch.cuda.mem_get_info()
    held0 = torch.cuda.memory_allocated()
    _lib.check(lib.p2p_instnorm_stats_static(_p(y), _p(got), N, H * W, C, eps, _DT[y.dtype], _stream()))
    torch.cuda.synchronize()
    free1, _ = torch.cuda.mem_get_info()
    return want, got, (free0, free1, held0, torch.cuda.memory_allocated())


def _check(want, got):
    assert torch.isfinite(got).all()
    assert relerr(got[..., 0].cpu().numpy(), want[..., 0].cpu().numpy()) < TOL      # mean
    assert relerr(got[..., 1].cpu().numpy(), want[..., 1].cpu().numpy()) < TOL      # 1 / sqrt(var + eps)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_two_calls_in_a_row_match_the_existing_entry_point(dtype):
    torch.manual_seed(11)
    y1 = (1.5 * torch.randn(8, 64, 33, 96, device="cuda") + 0.3).to(dtype)          # many row chunks per (image, channel block)
    y2 = (0.2 * torch.randn(3, 17, 9, 132, device="cuda") - 2.0).to(dtype)          # smaller, ragged channel block
    want1, got1, _ = _both(y1)
    want2, got2, mem = _both(y2)
    _check(want1, got1)
    _check(want2, got2)
    # the same tensor again gives the same answer (the scratch the first call left behind does not leak in)
    _, again, _ = _both(y1)
    _check(want1, again)
    # and against the definition (torch, fp32): biased variance over H*W
    x = y1.float()
    mean = x.mean(dim=(1, 2))
    rstd = torch.rsqrt(x.var(dim=(1, 2), unbiased=False) + 1e-5)
    assert relerr(got1[..., 0].cpu().numpy(), mean.cpu().numpy()) < 10 * TOL
    assert relerr(got1[..., 1].cpu().numpy(), rstd.cpu().numpy()) < 10 * TOL
    free0, free1, held0, held1 = mem
    assert held1 == held0                       # nothing taken from torch's allocator
    assert free1 >= free0                       # nothing taken from the device either (second call: modules are loaded)


def test_sizes_beyond_the_built_in_scratch_are_rejected():
    lib = _lib.lib()
    y = torch.zeros(1, 1, 1, 65536 + 4, device="cuda")                               # N*C > 65536
    out = torch.zeros(1, 65536 + 4, 2, device="cuda")
    rc = lib.p2p_instnorm_stats_static(_p(y), _p(out), 1, 1, 65536 + 4, 1e-5, 0, _stream())
    assert rc == -3                                                                  # P2P_ENOTSUP
    msg = lib.p2p_last_error().decode()
    assert "exceeds the built-in scratch" in msg and "65536" in msg
    y = torch.zeros(2, 2, 2, 6, device="cuda")                                       # C % 4 != 0: not covered either
    with pytest.raises(NotImplementedError):
        _lib.check(lib.p2p_instnorm_stats_static(_p(y), _p(out), 2, 4, 6, 1e-5, 0, _stream()))
    assert lib.p2p_instnorm_stats_static(_p(y), _p(out), 0, 4, 8, 1e-5, 0, _stream()) == 0   # empty: nothing to do
