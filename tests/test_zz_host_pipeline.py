"""host_pipeline = 1: b200spmv_spmv_host uploads x range by range (the columns each unit range of rows references) on a
third stream so that H2D overlaps D2H.  Same y as the default path and as the oracle.

Green on a B200 since round 2 (profiles/r02a_pytest_new_paths.txt); the path stays opt-in (opts.host_pipeline = 0 by default):
measured, it does not help matrices whose every row range references the whole band (profiles/r02a_bench_n1_pipeline.json)."""
import os

import numpy as np
import pytest

import matrices as M

pytestmark = [pytest.mark.gpu]


def _banded(m, per_row, half_band, seed):
    """m x m, every row has per_row entries inside [i - half_band, i + half_band] (clipped): each quarter of the rows
    references its own part of x, which is what the pipeline exploits."""
    rng = np.random.default_rng(seed)
    rows = np.repeat(np.arange(m, dtype=np.int64), per_row)
    cols = np.clip(rows + rng.integers(-half_band, half_band + 1, rows.size), 0, m - 1)
    import scipy.sparse as sp
    A = sp.coo_matrix((rng.uniform(-1, 1, rows.size), (rows, cols)), shape=(m, m)).tocsr()
    A.sort_indices()
    return A.indptr.astype(np.int32), A.indices.astype(np.int32), A.data.astype(np.float64)


def _random(m, per_row, seed):
    rng = np.random.default_rng(seed)
    import scipy.sparse as sp
    rows = np.repeat(np.arange(m, dtype=np.int64), per_row)
    A = sp.coo_matrix((rng.uniform(-1, 1, rows.size), (rows, rng.integers(0, m, rows.size))), shape=(m, m)).tocsr()
    A.sort_indices()
    return A.indptr.astype(np.int32), A.indices.astype(np.int32), A.data.astype(np.float64)


@pytest.mark.parametrize("fmt_name", ["csr", "sell"])
def test_host_pipeline_same_y(native, oracle, fmt_name):
    import spmv_research_b200 as S
    fmt = {"csr": S.CSR_ADAPTIVE, "sell": S.SELL_C_SIGMA}[fmt_name]
    m = 600000
    rp, col, val = _banded(m, 6, 2000, 1)
    # "reversed band": row i references columns near m - 1 - i, so the ranges of x arrive right to left
    # (columns are then descending inside a row, which the product does not mind)
    cases = [("banded", (rp, col, val)), ("reversed band", (rp, (m - 1 - col).astype(np.int32), val)), ("random", _random(m, 5, 3))]
    for tag, (rp, col, val) in cases:
        nnz = int(rp[-1])
        ref = S.csr_to_format(rp, col, val, m, m, nnz, format=fmt, pin_host_vectors=1)
        MF = S.csr_to_format(rp, col, val, m, m, nnz, format=fmt, pin_host_vectors=1, host_pipeline=1)
        for seed in (5, 6):
            x = M.rand_x(m, seed)
            y0, y1 = np.full(m, 1.0), np.full(m, 2.0)
            ref.spmv(x, y0)
            MF.spmv(x, y1)
            assert np.array_equal(y0, y1), tag              # same kernels, same launches: bit-identical
            y_ref = oracle.csr_spmv(rp, col, val, x)
            assert oracle.within_bound(y1, y_ref, oracle.row_abs_bound(rp, col, val, x)), tag
        # entries of x outside the referenced band never travel: NaN there stays invisible
        lo, hi = MF.handle.col_range()
        xn = M.rand_x(m, 7)
        y_ref = oracle.csr_spmv(rp, col, val, xn)
        xn2 = xn.copy()
        xn2[:lo] = np.nan
        xn2[hi:] = np.nan
        if fmt == S.SELL_C_SIGMA:
            xn2[0] = xn[0]
        y = np.full(m, 3.0)
        MF.spmv(xn2, y)
        assert oracle.within_bound(y, y_ref, oracle.row_abs_bound(rp, col, val, xn)), tag
        ref.close()
        MF.close()
