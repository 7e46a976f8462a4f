"""GPU: b2d_groupnorm_fwd_static -- the allocation-free variant of b2d_groupnorm_fwd (scratch in a file-scope
`static __device__` array, include/b2d.h) -- called through the C-ABI and checked against the existing entry point on
two calls in a row (the second with other data and fewer samples, so stale scratch contents would show), against
torch's fp32 GroupNorm, and for its size limit.  No counterpart in the reference.
The file name sorts last on purpose."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

pytestmark = pytest.mark.gpu


def rel(a, b):
    a, b = a.detach().float().cpu(), b.detach().float().cpu()
    return ((a - b).norm() / (b.norm() + 1e-20)).item()


def _run_both(N, H, W, Cch, seed, act):
    from lipreading_video_generation_b200 import _lib as L, ops
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(N, H, W, Cch, generator=g) * 1.3 + 0.4).half().cuda()
    gamma = (1 + 0.1 * torch.randn(Cch, generator=g)).cuda()
    beta = (0.1 * torch.randn(Cch, generator=g)).cuda()
    y_ref, mean_ref, rstd_ref = ops.groupnorm_fwd(x, gamma, beta, act=act)        # existing entry point, caller workspace
    # outputs are the caller's; everything is allocated BEFORE the call under test
    y = torch.full_like(y_ref, float("nan"))
    mean = torch.full_like(mean_ref, float("nan"))
    rstd = torch.full_like(rstd_ref, float("nan"))
    a = ops._gn_args(x, None, gamma, beta, 1e-5, act, 0.0, 0)
    stream = torch.cuda.current_stream().cuda_stream
    torch.cuda.synchronize()
    before = torch.cuda.memory_stats()["allocation.all.allocated"]
    rc = L.lib().b2d_groupnorm_fwd_static(C.byref(a), y.data_ptr(), mean.data_ptr(), rstd.data_ptr(), stream)
    L.check(rc, "b2d_groupnorm_fwd_static")
    torch.cuda.synchronize()
    assert torch.cuda.memory_stats()["allocation.all.allocated"] == before        # nothing went through torch's allocator
    # same kernels, same arithmetic; the fp64 scratch is filled by atomics, so allow the tests' 16-bit-storage tolerance
    assert rel(y, y_ref) <= 2e-3
    assert torch.allclose(mean, mean_ref, atol=1e-5, rtol=1e-5)
    assert torch.allclose(rstd, rstd_ref, rtol=1e-4)
    xc = x.float().cpu().permute(0, 3, 1, 2)
    want = F.group_norm(xc, 32, gamma.cpu(), beta.cpu(), 1e-5)
    want = F.silu(want) if act else want
    assert rel(y.float().cpu().permute(0, 3, 1, 2), want) <= 2e-3


def test_static_scratch_groupnorm_equals_the_workspace_entry_point_on_two_calls_in_a_row():
    # 64x64x64 and 32x32x256 halves are 512 KiB per sample: beyond the one-cluster plan, so the two-kernel path that
    # needs the scratch runs; the second call reuses the array with fewer samples and other values
    _run_both(5, 64, 64, 64, seed=11, act=True)
    _run_both(2, 32, 32, 256, seed=12, act=False)
    _run_both(5, 64, 64, 64, seed=13, act=True)


def test_static_scratch_groupnorm_small_samples_take_the_cluster_path_and_still_match():
    _run_both(3, 8, 8, 128, seed=14, act=True)


def test_static_scratch_groupnorm_rejects_more_samples_than_the_built_in_scratch_holds():
    from lipreading_video_generation_b200 import _lib as L, ops
    x = torch.zeros(1, 8, 8, 64, device="cuda", dtype=torch.float16)
    gamma, beta = torch.ones(64, device="cuda"), torch.zeros(64, device="cuda")
    a = ops._gn_args(x, None, gamma, beta, 1e-5, True, 0.0, 0)
    a.N = 1025                                                                   # B2D_GN_STATIC_MAX_N + 1
    y = torch.empty_like(x)
    m = torch.empty(1, 32, device="cuda")
    rc = L.lib().b2d_groupnorm_fwd_static(C.byref(a), y.data_ptr(), m.data_ptr(), m.data_ptr(),
                                          torch.cuda.current_stream().cuda_stream)
    assert rc == 4                                                               # B2D_ERR_WORKSPACE
    msg = L.lib().b2d_last_error().decode()
    assert "B2D_GN_STATIC_MAX_N" in msg and "1025" in msg
    torch.cuda.synchronize()                                                     # nothing was launched, nothing faulted
