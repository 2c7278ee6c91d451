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
