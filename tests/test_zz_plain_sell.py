"""SURVEY 8a row a8 -- the reference's plain SELL (K/sell.cpp: slices of VL = 4 / 8 consecutive rows, no sorting).
Its layout is pinned in tests/test_oracle_pin.py::test_plain_sell_vs_reference.  On the B200 the same function is
SELL-C-sigma with C = sigma = VL: the same slices (sorting inside a slice only permutes its lanes, widths and members are
unchanged), the same per-row chain of fused multiply-adds in ascending k, so y is BIT-equal to the plain SELL product
(padding entries add 0 * x[col] = +0 to a finite sum whichever column they carry)."""
import os

import numpy as np
import pytest

import matrices as M


def _cases():
    rng = np.random.default_rng(5)
    return [("ragged", M.csr_from_lengths(rng.integers(0, 9, 1003), 500, 1), 500),
            ("empty head", M.csr_from_lengths(np.r_[np.zeros(9, np.int64), rng.integers(0, 40, 500)], 2000, 2), 2000),
            ("short", M.csr_from_lengths(rng.integers(1, 6, 20000), 30000, 3), 30000)]


def test_sellcs_with_sigma_equal_C_is_plain_sell(oracle):
    """CPU: the SELL-C-sigma oracle at C = sigma = VL gives the plain SELL oracle's y bit for bit, for any thread count of the
    plain SELL conversion (only padding columns move)."""
    for name, (rp, col, val), n in _cases():
        m = len(rp) - 1
        x = M.rand_x(n, 3)
        for VL in (4, 8):
            O = oracle.sellcs_convert(rp, col, val, VL, VL)
            y = oracle.sellcs_spmv(O, x)
            for nt in (1, 3):
                P = oracle.sell_plain_convert(rp, col, val, VL, nt)
                assert np.array_equal(P["row_cluster_ptr"][1:] - P["row_cluster_ptr"][:-1], np.asarray(O["slice_width"], np.int64) * VL)
                assert np.array_equal(y, oracle.sell_plain_spmv(P, x)[:m]), (name, VL, nt)


@pytest.mark.gpu
def test_gpu_sell_with_sigma_equal_C_is_plain_sell(native, oracle):
    import spmv_research_b200 as S
    for name, (rp, col, val), n in _cases():
        m = len(rp) - 1
        x = M.rand_x(n, 3)
        for VL in (4, 8):
            MF = S.csr_to_format(rp, col, val, m, n, int(rp[-1]), format=S.SELL_C_SIGMA, sell_C=VL, sell_sigma=VL)
            y = np.full(m, 1.0)
            MF.spmv(x, y)
            P = oracle.sell_plain_convert(rp, col, val, VL, 1)
            y_plain = oracle.sell_plain_spmv(P, x)[:m]
            assert oracle.within_bound(y, y_plain, oracle.row_abs_bound(rp, col, val, x)), (name, VL)
            assert np.array_equal(y, y_plain), (name, VL)          # same slices, same per-row FMA chain: bit for bit (green on a B200, round 2)
            MF.close()
