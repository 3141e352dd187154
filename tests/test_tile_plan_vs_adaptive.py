"""the reference's CSR-Adaptive row blocking (K/GPU_clean/csr_adaptive_cuda.cu:46-79, restated as
oracle_adaptive_rowblocks) is the behavioural reference of the product's binning.  The product does not block by rows: it
cuts the nnz stream into windows of T non-zeros (csr_plan_tiles; the numpy model below is the one
tests/test_gpu_parity.py::test_csr_split_rows_and_tile_plan pins the device plan on, entry for entry).  This test ties
the two on random and edge-case row-length profiles, with BLOCK_SIZE = T:

 * both cover [0, m) with consecutive, non-overlapping row ranges;
 * a multi-row adaptive block (<= T non-zeros by construction) lies inside at most two consecutive windows: the work one
   CSR-Adaptive CTA reduces in shared memory is what at most two tile CTAs see;
 * a row CSR-Adaptive gives a block of its own because it is longer than T (its 'CSR-vector / long row' bin) always leaves its
   owner window: it is among the rows the tile plan finishes by tail extension or by the split-row fold;
 * the tile plan is the balanced one: ceil(nnz / T) units of exactly T non-zeros, never more units than the row blocking.
"""
import numpy as np
import pytest

import matrices as M


def tile_plan_model(rp, T, ext):
    """first row starting in each window, rows leaving their owner window (by <= ext: extension, by more: split)."""
    m, nnz = len(rp) - 1, int(rp[-1])
    ntiles = max(1, -(-nnz // T))
    first = np.minimum(np.searchsorted(rp, np.arange(ntiles) * T, side="left"), m).astype(np.int64)
    first[0] = 0
    first = np.append(first, m)
    leaving = {}
    for t in range(ntiles):
        r0, r1 = first[t], first[t + 1]
        over = int(rp[r1]) - (t + 1) * T
        if r1 > r0 and over > 0:
            leaving[int(r1 - 1)] = "split" if over > ext else "ext"
    return first, leaving


def _profiles():
    rng = np.random.default_rng(3)
    out = [(name, rp) for name, rp, _, _, _ in M.edge_cases()]
    out.append(("uniform10", np.concatenate([[0], np.cumsum(rng.integers(5, 16, 30000))]).astype(np.int32)))
    out.append(("skewed", np.concatenate([[0], np.cumsum(np.where(rng.random(8000) < 0.01, rng.integers(3000, 9000, 8000), rng.integers(0, 20, 8000)))]).astype(np.int32)))
    out.append(("long500", np.concatenate([[0], np.cumsum(rng.integers(300, 700, 900))]).astype(np.int32)))
    return out


@pytest.mark.parametrize("T,ext", [(2048, 256), (1024, 64), (4096, 2048)])
def test_tile_plan_against_csr_adaptive_rowblocks(oracle, T, ext):
    for name, rp in _profiles():
        rp = np.ascontiguousarray(rp, np.int32)
        m, nnz = len(rp) - 1, int(rp[-1])
        if m == 0:
            continue
        first, leaving = tile_plan_model(rp, T, ext)
        rb = oracle.adaptive_rowblocks(rp, T, 1).astype(np.int64)
        # both are consecutive covers of [0, m)
        assert rb[0] == 0 and rb[-1] == m and np.all(np.diff(rb) >= 0), name
        assert first[0] == 0 and first[-1] == m and np.all(np.diff(first) >= 0), name
        blocks = [(int(a), int(b)) for a, b in zip(rb[:-1], rb[1:]) if b > a]
        assert sum(b - a for a, b in blocks) == m, name
        n_units_blocks = 0
        for a, b in blocks:
            w = int(rp[b]) - int(rp[a])
            if b - a > 1 and b == m and w > T:
                # reference quirk: the greedy loop runs i < m, so the LAST row's length is never added and the closing block can
                # exceed BLOCK_SIZE (csr_adaptive_cuda.cu:52: `for (i = 1; i < m; i++)`) -- nothing to tie the tile plan to
                n_units_blocks += -(-w // T)
            elif b - a > 1:
                assert w <= T, (name, a, b, w)                      # multi-row block: fits one CTA's shared memory
                if w > 0:
                    t_lo, t_hi = int(rp[a]) // T, (int(rp[b]) - 1) // T
                    assert t_hi - t_lo <= 1, (name, a, b)           # ... and at most two consecutive windows of the tile plan
                n_units_blocks += 1
            else:
                if w > T:
                    # CSR-Adaptive's long-row bin: in the tile plan the row leaves its owner window (extension or split-row fold)
                    assert a in leaving, (name, a, w)
                n_units_blocks += max(1, -(-w // T))                # a long row costs its owner ceil(w / T) windows of work
        assert max(1, -(-nnz // T)) <= max(1, n_units_blocks), name
