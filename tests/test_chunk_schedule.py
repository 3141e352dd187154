"""b200spmv_chunk_schedule (csrc/power.cu): the row-chunk x column-panel launch order of the multi-GPU layer, a pure host
function of the panels' tile plans -- checked here without a GPU on random plans.

Invariants: every panel's tiles are covered exactly once, in order; the rows a panel finishes in chunk c were finished by
every earlier panel in chunks <= c (so its read-modify-write of y never meets an unwritten partial); final_row[c] is what
the LAST panel has finished before chunk c, non-decreasing, 0 at c = 0 and m at c = K."""
import numpy as np
import pytest


def _random_tile_rows(rng, m, ntiles):
    """first row starting in each tile: non-decreasing, [0] = 0, [ntiles] = m (repeats = tiles inside long rows)."""
    inner = np.sort(rng.integers(0, m + 1, size=ntiles - 1)) if ntiles > 1 else np.zeros(0, np.int64)
    return np.concatenate([[0], inner, [m]]).astype(np.int32)


@pytest.mark.parametrize("P,K", [(1, 1), (1, 4), (2, 4), (2, 8), (3, 5), (8, 16)])
def test_chunk_schedule_invariants(native, P, K):
    from spmv_research_b200 import capi
    rng = np.random.default_rng(100 * P + K)
    for trial in range(40):
        m = int(rng.integers(1, 5000))
        rows = [_random_tile_rows(rng, m, int(rng.integers(1, 200))) for _ in range(P)]
        unit, fin = capi.chunk_schedule(rows, K, m)
        assert fin[0] == 0 and fin[K] == m and np.all(np.diff(fin) >= 0)
        for q in range(P):
            nt = rows[q].size - 1
            assert unit[q, 0] == 0 and unit[q, K] == nt and np.all(np.diff(unit[q]) >= 0)
        for c in range(K + 1):
            done = [int(rows[q][unit[q, c]]) for q in range(P)]        # rows [0, done[q]) finished by panel q before chunk c
            if c == K:
                assert done == [m] * P
            for q in range(1, P):
                assert done[q] <= done[q - 1], (c, q, done)
            assert fin[c] == done[-1]


def test_chunk_schedule_rejects_bad_plans(native):
    from spmv_research_b200 import capi
    with pytest.raises(capi.B200Error):
        capi.chunk_schedule([np.array([1, 5], np.int32)], 2, 5)          # does not start at 0
    with pytest.raises(capi.B200Error):
        capi.chunk_schedule([np.array([0, 4], np.int32)], 2, 5)          # does not end at m
    with pytest.raises(capi.B200Error):
        capi.chunk_schedule([np.array([0, 5], np.int32)], 17, 5)         # too many chunks
