"""CPU tests of the drop-in boundary: libsvdj.so loads without a GPU, exports every symbol that
include/svdj.h declares, validates arguments before touching CUDA, and its host-side schedule is the
reference's (via the oracle).  No compute calls here."""
import ctypes as C
import os
import re

import numpy as np
import pytest

import helpers as H

ROOT = H.ROOT


def _declared_functions(header):
    txt = open(header).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b(svdj_[a-z0-9_]+)\s*\(", txt)))


def test_library_exports_every_declared_symbol(svdj):
    declared = _declared_functions(os.path.join(ROOT, "include", "svdj.h"))
    assert declared, "header parse failed"
    for name in declared:
        assert hasattr(svdj.lib, name), f"{name} declared in include/svdj.h but not exported"
    assert sorted(svdj.EXPORTS) == declared


def test_version_and_defaults(svdj):
    assert svdj.lib.svdj_version() >= 100
    o = svdj.default_opts()
    assert o.sort_desc == 1 and o.max_sweeps == 60 and o.threshold_mode == svdj.THRESH_REL and o.device == -1


@pytest.mark.parametrize("n", [2, 5, 16, 63, 64, 257])
def test_schedule_matches_oracle(svdj, n):
    assert svdj.num_steps(n) == H.oracle_lib().oracle_num_steps(n)
    for st in range(svdj.num_steps(n)):
        assert np.array_equal(svdj.schedule_step(n, st), H.oracle_schedule(n, st))
    assert len(svdj.schedule_step(n, svdj.num_steps(n))) == 0  # past the last step


@pytest.mark.parametrize("nb,vranks", [(4, 1), (4, 2), (8, 2), (16, 4), (24, 3), (64, 8), (128, 4), (128, 2), (256, 4), (12, 2), (9, 0)])
def test_block_schedule_visits_every_block_pair_once(svdj, nb, vranks):
    """The block path's orderings (flat round-robin and the two-level tournament): disjoint pairs in every step,
    every unordered block pair exactly once per sweep -- the property the reference's ordering has for columns
    (main.cu:498-545)."""
    seen = set()
    steps = svdj.num_steps(nb)
    for k in range(steps):
        pq = svdj.block_schedule_step(nb, vranks, k)
        assert len(pq) == nb // 2
        used = set()
        for p, q in pq.tolist():
            if p >= nb or q >= nb:   # odd nb: the idle block of the step is paired with the dummy index nb
                continue
            assert p != q and p not in used and q not in used
            used |= {p, q}
            key = (min(p, q), max(p, q))
            assert key not in seen
            seen.add(key)
            assert p < q                 # lower block first: the eigenvalue ordering needs a consistent direction
    assert len(seen) == nb * (nb - 1) // 2


def test_block_schedule_two_level_structure(svdj):
    """Round 0 stays inside super-block pairs (r, P + r); later rounds only cross the two super-blocks one
    virtual rank holds, and within a round every virtual rank keeps the same two super-blocks (what lets the
    multi-GPU ring keep them resident for the round)."""
    nb, P = 64, 4
    wb = nb // (2 * P)
    for k in range(2 * wb - 1):
        for p, q in svdj.block_schedule_step(nb, P, k).tolist():
            assert {p // wb, q // wb} <= {(p // wb) % P, (p // wb) % P + P}
    k = 2 * wb - 1
    while k < nb - 1:
        held = None
        for t in range(wb):
            pq = svdj.block_schedule_step(nb, P, k + t)
            sb = [tuple(sorted({p // wb, q // wb})) for p, q in pq.tolist()]
            assert all(len(x) == 2 for x in sb)       # cross pairs only
            by_rank = [sb[r * wb] for r in range(P)]
            assert all(sb[r * wb + i] == by_rank[r] for r in range(P) for i in range(wb))
            held = held or by_rank
            assert by_rank == held
        k += wb


def test_block_schedule_rejects_bad_arguments(svdj):
    assert len(svdj.block_schedule_step(10, 4, 0)) == 0     # 10 blocks do not split into 8 super-blocks
    assert len(svdj.block_schedule_step(16, 4, 15)) == 0    # step out of range
    assert len(svdj.block_schedule_step(1, 0, 0)) == 0


def test_argument_errors_come_back_as_negative_argument_numbers(svdj):
    A = np.zeros((4, 4), order="F")
    s = np.zeros(4)
    V = np.zeros((4, 4), order="F")
    call = svdj.lib.svdj_dgesvd
    pa, ps, pv = A.ctypes.data, s.ctypes.data, V.ctypes.data
    assert call(9, 0, 4, 4, pa, 4, ps, pv, 4, None, None) == -1
    assert call(0, 3, 4, 4, pa, 4, ps, pv, 4, None, None) == -2
    assert call(0, 0, 4, 4, None, 4, ps, pv, 4, None, None) == -5
    assert call(0, 0, 4, 4, pa, 3, ps, pv, 4, None, None) == -6
    assert call(0, 0, 4, 4, pa, 4, None, pv, 4, None, None) == -7
    assert call(0, 0, 4, 4, pa, 4, ps, None, 4, None, None) == -8
    assert call(0, 0, 4, 4, pa, 4, ps, pv, 3, None, None) == -9
    assert "argument" in svdj.last_error()
    with pytest.raises(svdj.SvdjError):
        svdj.dgesvd(0, 0, 4, 4, A, 2, s, V, 4)


def test_python_mirror_rejects_row_major_input(svdj):
    A = np.zeros((4, 6))  # C order
    with pytest.raises(ValueError):
        svdj.dgesvd(0, 0, 4, 6, A, 4, np.zeros(4), np.zeros((6, 6), order="F"), 6)


def test_no_cpu_fallback_without_a_device(svdj):
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    A = np.asfortranarray(np.eye(4))
    with pytest.raises(svdj.SvdjError) as e:
        svdj.dgesvd(0, 0, 4, 4, A, 4, np.zeros(4), np.zeros((4, 4), order="F"), 4)
    assert e.value.code in (svdj.ERR_CUDA, svdj.ERR_ALLOC)


def test_rows_entry_point_arguments_and_no_fallback(svdj):
    """svdj_dgesvd_rows_device: bad arguments come back as negative argument numbers before anything touches the
    device; with valid arguments and no GPU it fails with a CUDA error -- it never computes on the host."""
    import ctypes as C
    import torch
    cb = svdj.AllreduceFn(lambda ctx, ptr, count, stream: 0)
    st = svdj.Stats()
    p = lambda v: C.c_void_p(v)   # never dereferenced on the host

    def call(m_local, n, vrows, v_row0, opts, fn=cb):
        return svdj.lib.svdj_dgesvd_rows_device(svdj.AllVec, svdj.AllVec, m_local, n, p(0x1000), max(m_local, 1), p(0x2000),
                                                p(0x3000), max(vrows, 1), vrows, v_row0, C.byref(opts), C.byref(st), None,
                                                fn, None)

    with_tol = svdj.default_opts(tol=1e-12)
    assert call(0, 256, 256, 0, with_tol) == -3                     # no rows
    assert call(1000, 64, 64, 0, with_tol) == -4                    # n < 128: no 64-column block pairs
    assert call(1000, 256, 200, 100, with_tol) == -10               # V rows 100..299 of a 256-row V
    assert call(1000, 256, 256, 0, svdj.default_opts()) == -12      # tol must be set: the local m is not the global one
    assert b"tol" in svdj.lib.svdj_last_error()
    assert call(1000, 256, 256, 0, with_tol, svdj.AllreduceFn(0)) == -15   # no all-reduce
    if not torch.cuda.is_available():
        assert call(1000, 256, 256, 0, with_tol) in (svdj.ERR_CUDA, svdj.ERR_ALLOC)
