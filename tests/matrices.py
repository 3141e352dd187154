"""Seeded test inputs: generator parameter lines and hand-made CSR edge cases."""
import numpy as np

# small lines in the style of matrix_generation_parameters/*.txt (seed 14 as in the reference files)
GOLDEN_LINES = [
    "17099 17099 5.6078784547 4.3921621102 normal random 0.2972525308 61.9471560146 0.8033653966 0.6330185674 14",  # P/test.txt (scircuit twin)
    "20000 20000 10 5 normal random 0.3 0 0.5 0.5 14",               # config-1 shape, scaled down
    "20000 20000 50 150 gamma random 0.6 0 0.05 0.05 14",            # G1 heavy tail (about half the rows empty)
    "20000 30000 50 16.6667 gamma random 0.6 1000 0.05 0.05 14",     # G2 giant-row head, rectangular
    "5000 5000 20 6 normal diagonal 0.05 0 1.9 0.95 14",
    "3000 3000 500 166.6667 normal random 0.6 100 1.4 0.95 14",      # long rows
    "4191 4191 500 166.6667 normal random 0.6 10000 1.4 0.95 14",    # last lines of the medium dataset
]


def rand_x(n, seed=14):
    """x ~ U[-1, 1): exercises the gather (the reference harness only uses x = 1)."""
    return np.random.default_rng(seed).uniform(-1.0, 1.0, n)


def csr_from_lengths(lengths, n, seed=0):
    """CSR with the given row lengths, sorted unique random columns, values U[0,1)."""
    rng = np.random.default_rng(seed)
    lengths = np.asarray(lengths, np.int64)
    m = len(lengths)
    row_ptr = np.zeros(m + 1, np.int32)
    np.cumsum(lengths, out=row_ptr[1:])
    nnz = int(row_ptr[-1])
    col = np.zeros(nnz, np.int32)
    for i in range(m):
        l = int(lengths[i])
        if l:
            if l > n:
                raise ValueError("row longer than n")
            col[row_ptr[i]:row_ptr[i + 1]] = np.sort(rng.choice(n, l, replace=False)) if l < n // 2 or n < 1 << 16 \
                else np.unique(rng.integers(0, n, 2 * l))[:l]
    val = rng.uniform(0.0, 1.0, nnz)
    return row_ptr, col, val


def edge_cases():
    """(name, row_ptr, col, val, n) covering what the reference harness probes: empty rows
    (bench_spmv.cpp:606-609), ragged tails, rows crossing / exceeding tile windows."""
    rng = np.random.default_rng(7)
    cases = []
    cases.append(("all_empty", *csr_from_lengths([0] * 300, 50), 50))
    cases.append(("single_entry", *csr_from_lengths([0, 0, 1, 0], 10), 10))
    cases.append(("one_row", *csr_from_lengths([37], 64), 64))
    cases.append(("m1_empty", *csr_from_lengths([0], 5), 5))
    cases.append(("short_rows", *csr_from_lengths(rng.integers(0, 12, 5000), 4000, 1), 4000))
    cases.append(("exact_tile", *csr_from_lengths([8] * 256 * 3, 1000, 2), 1000))          # nnz multiple of the 2048 tile
    cases.append(("trailing_empty_after_full_tile", *csr_from_lengths([8] * 256 + [0] * 17, 1000, 3), 1000))
    cases.append(("leading_empty", *csr_from_lengths([0] * 100 + [5] * 900, 800, 4), 800))
    cases.append(("mid_rows", *csr_from_lengths(rng.integers(200, 900, 300), 5000, 5), 5000))   # tail extension on every tile
    cases.append(("ext_boundary", *csr_from_lengths([2047, 1025, 1024, 1023, 3, 2048, 1, 1024 + 2048, 7], 8000, 6), 8000))
    cases.append(("giant_row_first", *csr_from_lengths([30000] + [5] * 2000, 40000, 8), 40000))
    cases.append(("giant_rows_mixed", *csr_from_lengths([3] * 50 + [9000] + [0] * 40 + [4100, 4097, 2] + [11] * 500 + [7000], 20000, 9), 20000))
    cases.append(("giant_row_last", *csr_from_lengths([4] * 1000 + [25000], 30000, 10), 30000))
    cases.append(("many_empty_runs", *csr_from_lengths(np.where(rng.random(20000) < 0.9, 0, rng.integers(1, 60, 20000)), 3000, 11), 3000))
    cases.append(("rect_wide", *csr_from_lengths(rng.integers(0, 30, 700), 100000, 12), 100000))
    cases.append(("rect_tall", *csr_from_lengths(rng.integers(0, 6, 9000), 17, 13), 17))
    return cases


def spd_matrix(n, per_row, seed, dominance=1.05):
    """Symmetric, strictly diagonally dominant CSR (int32 / f64, sorted columns): S = B + B^T off the diagonal with
    B random, diag = dominance * sum |off-diagonal| + 0.1 -- what the reference's CG harness expects (an SPD system)."""
    import scipy.sparse as sp
    rng = np.random.default_rng(seed)
    nnz = n * per_row
    B = sp.coo_matrix((rng.uniform(-1.0, 1.0, nnz), (rng.integers(0, n, nnz), rng.integers(0, n, nnz))), shape=(n, n)).tocsr()
    B.setdiag(0)
    S = (B + B.T).tocsr()
    S.eliminate_zeros()
    d = dominance * np.asarray(abs(S).sum(axis=1)).ravel() + 0.1
    S = (S + sp.diags(d)).tocsr()
    S.sort_indices()
    return S.indptr.astype(np.int32), S.indices.astype(np.int32), S.data.astype(np.float64)


def dd_matrix(n, per_row, seed, dominance=1.05):
    """NONsymmetric, strictly row-diagonally dominant CSR (int32 / f64, sorted columns, every diagonal entry present):
    the kind of system the reference's BiCGSTAB harness is for."""
    import scipy.sparse as sp
    rng = np.random.default_rng(seed)
    nnz = n * per_row
    B = sp.coo_matrix((rng.uniform(-1.0, 1.0, nnz), (rng.integers(0, n, nnz), rng.integers(0, n, nnz))), shape=(n, n)).tocsr()
    B.setdiag(0)
    B.eliminate_zeros()
    d = dominance * np.asarray(abs(B).sum(axis=1)).ravel() + 0.1
    S = (B + sp.diags(d)).tocsr()
    S.sort_indices()
    return S.indptr.astype(np.int32), S.indices.astype(np.int32), S.data.astype(np.float64)


def convdiff_matrix(nx, c=0.4):
    """2-D convection-diffusion 5-point stencil on an nx x nx grid (nonsymmetric for c != 0): diag 4, west -1-c, east -1+c,
    south -1-c, north -1+c.  BiCGSTAB needs a few hundred iterations on it, which exercises the every-100 checkpoints."""
    import scipy.sparse as sp
    n = nx * nx
    T = sp.diags([(-1.0 - c) * np.ones(nx - 1), 4.0 * np.ones(nx), (-1.0 + c) * np.ones(nx - 1)], [-1, 0, 1])
    E = sp.diags([(-1.0 - c) * np.ones(nx - 1), (-1.0 + c) * np.ones(nx - 1)], [-1, 1])
    S = (sp.kron(sp.identity(nx), T) + sp.kron(E, sp.identity(nx))).tocsr()
    S.sort_indices()
    assert S.shape == (n, n)
    return S.indptr.astype(np.int32), S.indices.astype(np.int32), S.data.astype(np.float64)
