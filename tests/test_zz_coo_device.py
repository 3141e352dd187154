"""b200spmv_coo_to_csr / b200spmv_create_coo: COO -> CSR on the GPU, bit-identical to the oracle's coo_to_csr
(pinned on the reference's import path), and .mtx file -> format handle -> product within the per-row bound.

Green on a B200 since round 2's first GPU pass (profiles/r02a_pytest_new_paths.txt); the kernel source is also replayed on the CPU
(tests/emul/)."""
import os

import numpy as np
import pytest

pytestmark = [pytest.mark.gpu]

CASES = [(1, 1, 0, False), (5, 5, 1, False), (300, 450, 4000, False), (50, 3000, 60000, False), (10, 100000, 300000, False),
         (2000, 2000, 30000, True), (3, 40, 100, True), (1000000, 1000000, 3000000, False), (64, 64, 33 * 64, False)]


def _coo(m, n, nnz, dup, seed=5):
    rng = np.random.default_rng(seed)
    if dup:
        row, col = rng.integers(0, m, nnz).astype(np.int32), rng.integers(0, n, nnz).astype(np.int32)
    else:
        p = rng.choice(m * n, nnz, replace=False) if nnz else np.zeros(0, np.int64)
        row, col = (p // n).astype(np.int32), (p % n).astype(np.int32)
    return row, col, rng.normal(size=nnz)


@pytest.mark.parametrize("m,n,nnz,dup", CASES)
def test_coo_to_csr_device_bit_identical(native, oracle, m, n, nnz, dup):
    import torch
    from spmv_research_b200 import capi
    row, col, val = _coo(m, n, nnz, dup)
    ref = oracle.coo_to_csr(row, col, val, m, n)
    dev = "cuda:0"
    d_r, d_c, d_v = (torch.from_numpy(a).to(dev) for a in (row, col, val))
    d_rp = torch.full((m + 1,), -1, dtype=torch.int32, device=dev)
    d_co = torch.full((max(nnz, 1),), -1, dtype=torch.int32, device=dev)
    d_vo = torch.full((max(nnz, 1),), -1.0, dtype=torch.float64, device=dev)
    dups = capi.coo_to_csr_device(d_r, d_c, d_v, m, n, nnz, d_rp, d_co, d_vo, torch.cuda.current_stream().cuda_stream)
    assert dups == ref[3]
    assert np.array_equal(d_rp.cpu().numpy(), ref[0])
    assert np.array_equal(d_co.cpu().numpy()[:nnz], ref[1]) and np.array_equal(d_vo.cpu().numpy()[:nnz], ref[2])


def test_coo_errors(native):
    import torch
    from spmv_research_b200 import capi
    import spmv_research_b200 as S
    r = torch.tensor([0, 5], dtype=torch.int32, device="cuda")
    c = torch.tensor([0, 0], dtype=torch.int32, device="cuda")
    v = torch.ones(2, dtype=torch.float64, device="cuda")
    out = torch.zeros(8, dtype=torch.int32, device="cuda")
    with pytest.raises(capi.B200Error, match="outside"):
        capi.coo_to_csr_device(r, c, v, 5, 5, 2, out, out, v.clone())
    with pytest.raises(capi.B200Error, match="duplicate"):
        S.coo_to_format(np.array([1, 1], np.int32), np.array([2, 2], np.int32), np.ones(2), 4, 4)


def test_mtx_file_to_product(native, oracle, tmp_path):
    """import_file() + csr_to_format() + spmv for Matrix Market files (general, symmetric, pattern, skew, Hermitian ...):
    the handle built from the file multiplies like the oracle's serial CSR kernel on the oracle's CSR of the same file."""
    from spmv_research_b200 import mtx, capi
    from test_mtx import _files
    for path in _files(str(tmp_path)):
        row, col, val, info = mtx.read_mtx(path)
        rp, ci, va, dups = oracle.coo_to_csr(row, col, val, info["m"], info["n"])
        assert dups == 0
        x = np.random.default_rng(14).uniform(-1, 1, info["n"])
        y_ref = oracle.csr_spmv(rp, ci, va, x)
        bound = oracle.row_abs_bound(rp, ci, va, x)
        for fmt in (capi.CSR_ADAPTIVE, capi.SELL_C_SIGMA):
            MF, info2 = mtx.mtx_to_format(path, format=fmt)
            assert (MF.m, MF.n, MF.nnz) == (info["m"], info["n"], info["nnz"])
            y = np.full(info["m"], 1.0)
            MF.spmv(x, y)
            assert np.all(np.abs(y - y_ref) <= 1e-12 * bound), path      # north-star per-row bound
            MF.close()
