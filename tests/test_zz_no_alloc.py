"""gsb_isotropic_reg_static: the allocation-free variant of gsb_isotropic_reg's deterministic mode (scratch in
a file-scope `static __device__` array of the library; include/gsb.h).  Called through the C ABI and checked
against the existing entry point, twice in a row with different sizes (the second call finds the first call's
partials in the shared array).  Named test_zz_* so that pytest collects it last."""
import pytest
import torch

from gslam_b200._lib import GsbError, call, lib, ptr, stream

pytestmark = pytest.mark.gpu


def _inputs(n, seed, dev):
    g = torch.Generator().manual_seed(seed)
    s = (-4.0 + 0.7 * torch.randn(n, 3, generator=g)).to(dev)
    vis = torch.randint(0, 3, (n,), generator=g).to(torch.int32).to(dev)
    grad0 = torch.randn(n, 3, generator=g).to(dev)
    return s, vis, grad0


def _existing(s, vis, grad0, weight, deterministic):
    n = s.shape[0]
    grad = grad0.clone()
    loss = torch.zeros(1, device=s.device)
    part = torch.empty((n + 255) // 256, device=s.device) if deterministic else None
    call("gsb_isotropic_reg", ptr(s), ptr(vis), n, weight, ptr(grad), ptr(loss), ptr(part), stream())
    return grad, loss


def _static(s, vis, grad0, weight, loss):
    grad = grad0.clone()
    call("gsb_isotropic_reg_static", ptr(s), ptr(vis), s.shape[0], weight, ptr(grad), ptr(loss), stream())
    return grad


def test_static_scratch_variant_matches_the_existing_entry_point_twice_in_a_row(cuda):
    weight = 5e-4
    loss = torch.full((1,), 123.0, device=cuda)          # stale value: the entry point overwrites it
    for n, seed in ((200_003, 0), (70_001, 1), (200_003, 0)):      # larger, smaller (stale partials behind it), again
        s, vis, grad0 = _inputs(n, seed, cuda)
        g_det, l_det = _existing(s, vis, grad0, weight, True)
        g_atm, l_atm = _existing(s, vis, grad0, weight, False)
        g_new = _static(s, vis, grad0, weight, loss)
        torch.cuda.synchronize()
        # same kernel, same fixed-order sum: bit-identical to the deterministic mode with caller scratch ...
        assert torch.equal(g_new, g_det) and torch.equal(loss, l_det)
        # ... and equal to the atomic mode within the tolerance the other tests use for this kernel
        assert torch.allclose(g_new, g_atm, rtol=1e-5, atol=1e-7)
        assert abs(float(loss) - float(l_atm)) <= 1e-4 * abs(float(l_atm))
        assert float(loss) > 0


def test_static_scratch_variant_limits_and_empty_input(cuda):
    L = lib()
    max_n = L.gsb_isotropic_reg_static_max_n()
    assert max_n == 256 * 65536
    s, vis, grad0 = _inputs(1000, 2, cuda)
    loss = torch.full((1,), 7.0, device=cuda)
    # beyond the maximum: rejected before anything is launched or read, with a clear message
    with pytest.raises(GsbError, match="exceeds the static scratch"):
        call("gsb_isotropic_reg_static", ptr(s), ptr(vis), max_n + 1, 5e-4, ptr(grad0), ptr(loss), stream())
    with pytest.raises(GsbError):      # loss_sum is required
        call("gsb_isotropic_reg_static", ptr(s), ptr(vis), 1000, 5e-4, ptr(grad0), None, stream())
    torch.cuda.synchronize()
    assert float(loss) == 7.0
    # N = 0: the sum of nothing
    call("gsb_isotropic_reg_static", None, None, 0, 5e-4, None, ptr(loss), stream())
    torch.cuda.synchronize()
    assert float(loss) == 0.0
    # at the maximum itself the call is accepted (the partial array is used to its last element)
    n = int(max_n)
    g = torch.Generator().manual_seed(3)
    s = (-4.0 + 0.5 * torch.randn(n, 3, generator=g)).to(cuda)
    vis = torch.ones(n, dtype=torch.int32, device=cuda)
    grad0 = torch.zeros(n, 3, device=cuda)
    g_det, l_det = _existing(s, vis, grad0, 5e-4, True)
    g_new = _static(s, vis, grad0, 5e-4, loss)
    torch.cuda.synchronize()
    assert torch.equal(g_new, g_det) and torch.equal(loss, l_det)
