"""fp32 (`_f` build) CSR product (csrc/f32.cu), checked WITHOUT a GPU: the kernel's tile bookkeeping
(csrc/f32_tile.h compiled as plain C++) replayed with the device's operation order (tests/emul/f32_emul.cpp -- test
infrastructure) writes every row exactly once and stays within the fp32 per-row bound of the oracle's `_f` serial kernel,
which is pinned bit for bit on the reference's `_f` build of K/csr.cpp (tests/test_oracle_pin.py).

fp32 tolerance (stated here, used by the GPU test too): |y - y_ref| <= 2 (len_i + 2) 2^-24 sum_j |a_ij x_j| per row -- both
sums are float accumulations of len_i terms, each within gamma_len of the exact sum."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import matrices as M
from spmv_research_b200 import amg

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "emul", "f32_emul.cpp")
SO = os.path.join(HERE, "emul", "libf32_emul.so")
HDR = os.path.join(HERE, "..", "spmv_research_b200", "csrc", "f32_tile.h")


def f32_tolerance(rp, bound):
    return 2.0 * (np.diff(rp) + 2) * 2.0 ** -24 * bound


def load_emul():
    if not os.path.exists(SO) or any(os.path.getmtime(d) > os.path.getmtime(SO) for d in (SRC, HDR)):
        subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-Wall", "-Wextra", "-o", SO, SRC])
    L = C.CDLL(SO)
    L.f32_emul_spmv.argtypes = [C.c_void_p] * 3 + [C.c_int64] * 3 + [C.c_void_p] * 3

    def run(rp, col, val, n, x):
        m = len(rp) - 1
        y = np.full(m, np.nan, np.float32)
        w = np.zeros(max(m, 1), np.int32)
        col = col if col.size else np.zeros(1, np.int32)
        val = val if val.size else np.zeros(1)
        L.f32_emul_spmv(rp.ctypes.data, col.ctypes.data, val.ctypes.data, m, n, int(rp[-1]), x.ctypes.data, y.ctypes.data, w.ctypes.data)
        return y, w[:m]
    return run


def edge_cases():
    rng = np.random.default_rng(3)
    out = []
    for name, lens, n in [("no non-zeros", [0] * 50, 10),
                          ("leading / trailing empty rows", [0] * 300 + [5] * 10 + [0] * 5000 + [3000] + [0] * 77, 5000),
                          ("one row over 5 windows", [0, 0, 9000, 0], 20000),
                          ("rows ending exactly on window ends", [2048] * 3 + [1024, 1024] + [0] * 3 + [2048], 4096),
                          ("mixed", list(rng.integers(0, 40, 3000)) + [5000] + list(rng.integers(0, 3, 4000)), 8000),
                          ("rows of 32", [32] * 640, 1000),
                          ("63 rows + a long one (warp bin)", [1] * 63 + [1985] + [7] * 100, 3000),
                          ("single entry", [0, 1, 0], 3)]:
        out.append((name,) + M.csr_from_lengths(np.array(lens), n, 7) + (n,))
    return out


@pytest.fixture(scope="module")
def emul():
    return load_emul()


def test_every_row_written_once_and_within_fp32_bound(emul, oracle, golden_matrices):
    cases = [(p.line(), rp, col, val, p.cols) for (p, rp, col, val) in golden_matrices] + edge_cases()
    rng = np.random.default_rng(3)
    for name, rp, col, val, n in cases:
        x = rng.uniform(-1, 1, n).astype(np.float32)
        y, w = emul(rp, col, val, n, x)
        assert np.all(w == 1), name
        y_ref = oracle.csr_spmv_f32(rp, col, val, x)
        tol = f32_tolerance(rp, oracle.row_abs_bound_f32(rp, col, val, x))
        assert np.all(np.abs(y.astype(np.float64) - y_ref.astype(np.float64)) <= tol), name
