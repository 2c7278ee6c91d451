"""svdj_step_device_noalloc: the split-path step for callers that cannot allocate device memory.

Some users of libsvdj drive the step loop from places where cudaMalloc is off the table: several
streams are in flight and must not be stalled by the allocator's device-wide synchronisation, or all
device memory already belongs to one arena, or each step has a latency budget.  For them the library
carries its own scratch array on every GPU.  This file checks, through the C ABI, that the variant gives
what the allocating entry point (svdj_step_device on its split path) gives, on two calls in a row that
share that array, and that problems too large for the array are refused with a clear message.

The file name sorts last on purpose: if this test fails under `pytest -x`, every other test has
already run.  It also passes alone in a fresh process (the first call does the per-GPU set-up).
"""
import numpy as np
import pytest

import helpers as H

torch = pytest.importorskip("torch")
pytestmark = pytest.mark.gpu


def _dev(a):
    return torch.from_numpy(np.ascontiguousarray(a.T)).cuda()


def _host(t, m, n):
    return np.asfortranarray(t.cpu().numpy().T[:m, :n])


@pytest.mark.parametrize("m,n", [(20000, 24), (3000, 64)])
def test_noalloc_step_matches_allocating_entry_point_twice_in_a_row(svdj, m, n):
    A0 = H.normal_matrix(m, n, seed=m)
    V0 = np.asfortranarray(np.eye(n))
    tol = m * 2.0 ** -53
    assert svdj.step_scratch_doubles(m, n) <= svdj.NOALLOC_MAX_SCRATCH_DOUBLES

    # existing entry point, split path, caller-provided scratch
    rA, rV = _dev(A0), _dev(V0)
    rctr = torch.zeros(2, dtype=torch.int64, device="cuda")
    scratch = torch.empty(svdj.step_scratch_doubles(m, n), dtype=torch.float64, device="cuda")
    # new entry point: nothing allocated by the call, scratch inside the library
    tA, tV = _dev(A0), _dev(V0)
    tctr = torch.zeros(2, dtype=torch.int64, device="cuda")
    torch.cuda.synchronize()

    cmp_tol = 64 * m * 2.0 ** -53  # the step tolerance of tests/test_l1_gpu.py
    for call, step in enumerate([0, 1]):  # two calls in a row: the second finds the first one's leftovers
        svdj.step_device(m, n, rA.data_ptr(), m, rV.data_ptr(), n, True, step, tol, path=svdj.PATH_SPLIT,
                         d_scratch=scratch.data_ptr(), d_counters=rctr.data_ptr())
        svdj.step_device_noalloc(m, n, tA.data_ptr(), m, tV.data_ptr(), n, True, step, tol,
                                 d_counters=tctr.data_ptr())
        torch.cuda.synchronize()
        ra, rv, ta, tv = _host(rA, m, n), _host(rV, n, n), _host(tA, m, n), _host(tV, n, n)
        assert np.abs(ta - ra).max() <= cmp_tol * np.abs(A0).max() * np.sqrt(m), f"A differs on call {call}"
        assert np.abs(tv - rv).max() <= cmp_tol, f"V differs on call {call}"
        assert int(tctr[0]) == int(rctr[0]) and int(tctr[0]) > 0
        assert int(tctr[1]) == int(rctr[1])
    assert not np.array_equal(_host(tA, m, n), A0), "the steps must have rotated something"


def test_noalloc_refuses_problems_larger_than_its_array(svdj):
    m, n = 8, 70000  # 35000 pairs x 1 split x 3 doubles > 98304
    assert svdj.step_scratch_doubles(m, n) > svdj.NOALLOC_MAX_SCRATCH_DOUBLES
    dummy = torch.zeros(16, dtype=torch.float64, device="cuda")
    with pytest.raises(svdj.SvdjError) as e:
        svdj.step_device_noalloc(m, n, dummy.data_ptr(), m, 0, n, False, 0, 1e-15)
    assert e.value.code == svdj.ERR_UNSUPPORTED
    assert "scratch" in str(e.value) and "98304" in str(e.value)
