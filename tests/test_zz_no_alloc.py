"""The allocation-free statistics entry point (pt_feature_stats_static) against the workspace one (pt_feature_stats).

Collected last on purpose (file name): it must pass alone and after every other test has used the library, because its
scratch is a process-wide array inside the library."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _both(x):
    from probtalk_b200 import _lib
    L = _lib.lib()
    st = torch.cuda.current_stream().cuda_stream
    N, D = x.shape
    nbytes = L.pt_feature_stats_workspace_bytes(D)
    ws = torch.empty(nbytes // 8, dtype=torch.float64, device="cuda")
    outs = []
    for static in (False, True):
        mean = torch.full((D,), float("nan"), dtype=torch.float64, device="cuda")
        cov = torch.full((D, D), float("nan"), dtype=torch.float64, device="cuda")
        if static:
            _lib.check(L.pt_feature_stats_static(x.data_ptr(), N, D, mean.data_ptr(), cov.data_ptr(), st))
        else:
            _lib.check(L.pt_feature_stats(x.data_ptr(), N, D, mean.data_ptr(), cov.data_ptr(), ws.data_ptr(), nbytes, st))
        outs.append((mean, cov))
    torch.cuda.synchronize()
    return outs


def test_static_scratch_matches_workspace_entry_point_twice_in_a_row():
    g = torch.Generator().manual_seed(5)
    # two calls in a row with different shapes: the second must not see anything of the first in the shared scratch
    for N, D in ((3000, 128), (517, 40)):
        x = (torch.randn(N, D, generator=g) * 1.5 + 0.25).cuda()
        (m0, c0), (m1, c1) = _both(x)
        assert torch.isfinite(m1).all() and torch.isfinite(c1).all()
        assert torch.equal(m0, m1) and torch.equal(c0, c1)      # same kernels, same reduction order: same bits
        ref = torch.cov(x.double().t().cpu())
        assert (c1.cpu() - ref).abs().max() < 1e-10


def test_static_scratch_rejects_oversize():
    from probtalk_b200 import _lib
    L = _lib.lib()
    x = torch.randn(16, 129).cuda()
    out = torch.empty(129 * 130, dtype=torch.float64, device="cuda")
    rc = L.pt_feature_stats_static(x.data_ptr(), 16, 129, out.data_ptr(), out.data_ptr() + 8 * 129,
                                   torch.cuda.current_stream().cuda_stream)
    assert rc != 0 and b"exceeds the built-in scratch" in L.pt_last_error()
