"""rs_convergence_test_noalloc: the allocation-free variant of convergence test 1.

Some users of this library run where device allocation is off limits once things are moving: an
application whose own pool owns the GPU heap, a coupled model stepping in real time where a
cudaMalloc would stall every stream on the device, a plugin host that forbids allocation after
start-up.  For them the library keeps a small static scratch array on the device and this entry
point works entirely out of it and the caller's own buffers.

The test drives it the way such a caller would -- device buffers created up front (torch is only the
allocator here), then two calls in a row on the same stream, reusing the shared scratch -- and checks
each result against the ordinary rs_convergence_test on the same numbers.  Tolerances are those of
test_gpu_parity.test_convergence_test_kernel: the max is exact, the sum may differ by summation order
(1e-13 relative).  It lives in a file of its own that pytest collects last, so a failure here cannot
hide the outcome of any other test; it must also pass when run alone in a fresh process.
"""
import numpy as np
import pytest

import helpers as H

pytestmark = pytest.mark.gpu


def _reference(lib, beta, ewt):
    from rootsolve_b200 import _capi
    out = np.zeros(2)
    _capi.check(lib.rs_convergence_test(len(beta), H.dp(beta), H.dp(ewt), H.dp(out)))
    return out


def test_noalloc_matches_existing_entry_point_twice_in_a_row(built):
    import torch
    from rootsolve_b200 import _capi
    rng = np.random.default_rng(20260101)
    nmax = 3_000_017
    d_beta = torch.empty(nmax, dtype=torch.float64, device="cuda")
    d_ewt = torch.empty(nmax, dtype=torch.float64, device="cuda")
    d_out = torch.zeros(2, dtype=torch.float64, device="cuda")
    torch.cuda.synchronize()
    # two back-to-back calls of different length: the second must not see anything of the first
    for n in (nmax, 70_001):
        beta = rng.standard_normal(n) * 1e-3
        ewt = rng.uniform(1e-9, 1e-6, n)
        d_beta[:n].copy_(torch.from_numpy(beta))
        d_ewt[:n].copy_(torch.from_numpy(ewt))
        torch.cuda.synchronize()
        _capi.check(built.rs_convergence_test_noalloc(n, d_beta.data_ptr(), d_ewt.data_ptr(), d_out.data_ptr(), None))
        torch.cuda.synchronize()
        got = d_out.cpu().numpy()
        want = _reference(built, beta, ewt)
        assert abs(got[0] - want[0]) <= 1e-13 * abs(want[0])
        assert got[1] == want[1]
        # and, independently of the library's other entry point, against numpy
        assert abs(got[0] - np.abs(beta).sum()) <= 1e-13 * np.abs(beta).sum()
        assert got[1] == np.max(np.abs(beta / ewt))


def test_noalloc_rejects_sizes_beyond_the_static_scratch(built):
    import torch
    from rootsolve_b200 import _capi
    limit = 8355840
    d = torch.zeros(8, dtype=torch.float64, device="cuda")
    rc = built.rs_convergence_test_noalloc(limit + 1, d.data_ptr(), d.data_ptr(), d.data_ptr(), None)
    assert rc == -10  # RS_ETOOBIG
    assert "exceeds the fixed-scratch limit" in built.rs_last_error().decode()
    with pytest.raises(_capi.RootSolveError):
        _capi.check(built.rs_convergence_test_noalloc(0, d.data_ptr(), d.data_ptr(), d.data_ptr(), None))
