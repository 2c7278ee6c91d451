"""CPU tests of the in-library multi-GPU driver's host logic (csrc/svdj_mg.cu) through svdj_mg_plan: the
tournament of the 2P column blocks and the four-buffer role tables that let a sender know which buffer of its
neighbour is free.  They stand where the reference's per-step rank assignment stood (main.cu:526-580)."""
import itertools

import pytest


@pytest.mark.parametrize("P", [1, 2, 3, 4, 8])
def test_every_block_pair_meets_exactly_once_per_sweep(svdj, P):
    seen = set()
    for rd in range(2 * P - 1):
        top, bottom, tbuf, bbuf, tr = svdj.mg_plan(P, rd)
        assert sorted(top + bottom) == list(range(2 * P))          # a permutation of the blocks
        for r in range(P):
            pair = frozenset((top[r], bottom[r]))
            assert pair not in seen
            seen.add(pair)
    assert len(seen) == P * (2 * P - 1)                             # all (2P choose 2) pairs
    top, bottom, *_ = svdj.mg_plan(P, 2 * P - 1)
    assert top == list(range(P)) and bottom == list(range(P, 2 * P))  # a sweep brings every block home


@pytest.mark.parametrize("P", [2, 3, 4, 8])
def test_transfers_realise_the_move_and_never_hit_a_live_buffer(svdj, P):
    for rd in range(3 * (2 * P - 1)):
        top, bottom, tbuf, bbuf, tr = svdj.mg_plan(P, rd)
        ntop, nbottom, ntbuf, nbbuf, _ = svdj.mg_plan(P, rd + 1)
        where = {}                                                   # (rank, buffer) -> block id before the move
        for r in range(P):
            where[(r, tbuf[r])] = top[r]
            where[(r, bbuf[r])] = bottom[r]
            assert tbuf[r] != bbuf[r] and 0 <= tbuf[r] < 4 and 0 <= bbuf[r] < 4
        after = dict(where)
        for (sr, sb, dr, db) in tr:
            assert (dr, db) not in where, "a transfer writes into a buffer its owner is still reading"
            assert abs(sr - dr) == 1, "blocks only travel between ring neighbours"
            after[(dr, db)] = where[(sr, sb)]
        for r in range(P):
            assert after[(r, ntbuf[r])] == ntop[r] and after[(r, nbbuf[r])] == nbottom[r]
        # every adjacent pair of ranks exchanges in both directions every round (the flags rely on it)
        dirs = {(sr, dr) for (sr, sb, dr, db) in tr}
        for r in range(P - 1):
            assert (r, r + 1) in dirs and (r + 1, r) in dirs
        # a rank sends at most one block to, and receives at most one from, each neighbour
        assert len(dirs) == len(tr)


def test_plan_rejects_bad_arguments(svdj):
    import ctypes as C
    assert svdj.lib.svdj_mg_plan(0, 0, None, None, None, None, None, None) < 0
    assert svdj.lib.svdj_mg_plan(2, -1, None, None, None, None, None, None) < 0
