"""b200spmv_amg_fill_device: pass 2 of the synthetic generator on the GPU, bit-identical to the host generator
(and so to the reference), whole matrices and row shards; the generated arrays feed b200spmv_create_device directly.

Green on a B200 since round 2's first GPU pass (profiles/r02a_pytest_new_paths.txt); the kernel source is also replayed on the CPU
(tests/emul/)."""
import os

import numpy as np
import pytest

import matrices as M
from spmv_research_b200 import amg

pytestmark = [pytest.mark.gpu]

LINES = list(M.GOLDEN_LINES) + [
    "5000 5000 20 6 normal simple 0.05 0 1.9 0.95 3",
    "2500 4000 8 30 gamma diagonal 0.9 3 0.0 0.0 99",
    "200000 200000 50 16.6667 gamma random 0.6 10000 0.05 0.05 14",
    "1000000 1000000 10 5 normal random 0.3 0 0.5 0.5 14",
]


@pytest.mark.parametrize("line", LINES)
def test_fill_device_bit_identical(native, line):
    import torch
    p = amg.MatrixParams.parse(line)
    rp, col, val = amg.generate(p)
    rp2, d_rp, d_col, d_val = amg.generate_device(p)
    assert np.array_equal(rp, rp2)
    assert np.array_equal(d_col.cpu().numpy(), col) and np.array_equal(d_val.cpu().numpy(), val)
    for a, b in [(1, p.rows // 3 + 5), (p.rows // 3 + 5, p.rows // 2), (p.rows // 2, p.rows - 1), (p.rows - 1, p.rows)]:
        _, _, c, v = amg.generate_device(p, a, b)
        assert np.array_equal(c.cpu().numpy(), col[rp[a]:rp[b]]) and np.array_equal(v.cpu().numpy(), val[rp[a]:rp[b]]), (a, b)
    torch.cuda.synchronize()


def test_generated_in_hbm_feeds_create_device(native, oracle):
    """host pass 1 -> device pass 2 -> b200spmv_create_device -> product, never touching the host with the non-zeros."""
    import torch
    from spmv_research_b200 import capi
    p = amg.MatrixParams.parse("200000 200000 10 5 normal random 0.3 100 0.5 0.5 14")
    rp, col, val = amg.generate(p)
    _, d_rp, d_col, d_val = amg.generate_device(p)
    x = np.random.default_rng(14).uniform(-1.0, 1.0, p.cols)
    y_ref = oracle.csr_spmv(rp, col, val, x)
    bound = oracle.row_abs_bound(rp, col, val, x)
    for fmt in (capi.CSR_ADAPTIVE, capi.SELL_C_SIGMA):
        h = capi.Handle(d_rp, d_col, d_val, p.rows, p.cols, int(rp[-1]), capi.default_opts(format=fmt), device_arrays=True)
        dx = torch.from_numpy(x).cuda()
        dy = torch.full((p.rows,), 1.0, dtype=torch.float64, device="cuda")
        h.spmv_device(dx, dy, torch.cuda.current_stream().cuda_stream)
        torch.cuda.synchronize()
        assert np.all(np.abs(dy.cpu().numpy() - y_ref) <= 1e-12 * bound)     # north-star per-row bound
        h.close()


def test_fill_device_errors(native):
    import torch
    from spmv_research_b200 import capi
    p = amg.MatrixParams(50, 50, 5, 1)
    d = torch.zeros(64, dtype=torch.int32, device="cuda")
    with pytest.raises(capi.B200Error, match="nr_rows >= 100"):
        amg.fill_device(p, d, 1.0, 0, 50, d, d)
    p = amg.MatrixParams.parse("1000 1000 5 1 normal random 0.3 0 0.5 0.5 14")
    with pytest.raises(capi.B200Error, match="bad row range"):
        amg.fill_device(p, d, 1.0, 5, 2000, d, d)
