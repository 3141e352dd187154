"""GPU COO -> CSR (csrc/coo.cu), checked WITHOUT a GPU: the conversion restated step for step around the very
sort networks the kernel uses (csrc/coo_sort.h compiled as plain C++ with 32 emulated lanes, tests/emul/coo_emul.cpp -- test
infrastructure), with the bucket scatter done in scrambled orders (the device's atomics give an arbitrary one), must equal the
oracle's coo_to_csr -- which is pinned on the reference's own import path -- bit for bit, duplicates included."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "emul", "coo_emul.cpp")
SO = os.path.join(HERE, "emul", "libcoo_emul.so")
HDR = os.path.join(HERE, "..", "spmv_research_b200", "csrc", "coo_sort.h")


@pytest.fixture(scope="module")
def emul():
    if not os.path.exists(SO) or any(os.path.getmtime(d) > os.path.getmtime(SO) for d in (SRC, HDR)):
        subprocess.check_call(["g++", "-O2", "-fPIC", "-shared", "-Wall", "-Wextra", "-o", SO, SRC])
    L = C.CDLL(SO)
    L.coo_emul_to_csr.restype = C.c_int64
    L.coo_emul_to_csr.argtypes = [C.c_void_p] * 3 + [C.c_int64] * 3 + [C.c_void_p] * 3 + [C.c_uint32]

    def run(row, col, val, m, n, scramble):
        nnz = len(row)
        rp, co, vo = np.zeros(m + 1, np.int32), np.zeros(max(nnz, 1), np.int32), np.zeros(max(nnz, 1))
        d = L.coo_emul_to_csr(row.ctypes.data, col.ctypes.data, val.ctypes.data, m, n, nnz, rp.ctypes.data, co.ctypes.data,
                              vo.ctypes.data, scramble)
        return rp, co[:nnz], vo[:nnz], d
    return run


CASES = [(1, 1, 0, False), (5, 5, 1, False), (300, 450, 4000, False), (50, 3000, 60000, False), (10, 100000, 300000, False),
         (2000, 2000, 30000, True), (3, 40, 100, True), (100000, 100000, 200000, False), (64, 64, 33 * 64, False)]


@pytest.mark.parametrize("m,n,nnz,dup", CASES)
def test_conversion_equals_oracle_for_any_scatter_order(emul, oracle, m, n, nnz, dup):
    rng = np.random.default_rng(5)
    if dup:
        row, col = rng.integers(0, m, nnz).astype(np.int32), rng.integers(0, n, nnz).astype(np.int32)
    else:
        p = rng.choice(m * n, nnz, replace=False) if nnz else np.zeros(0, np.int64)
        row, col = (p // n).astype(np.int32), (p % n).astype(np.int32)
    val = rng.normal(size=nnz)
    ref = oracle.coo_to_csr(row, col, val, m, n)
    assert (ref[3] > 0) == dup
    for scramble in (0, 1, 12345, 987654321):
        got = emul(row, col, val, m, n, scramble)
        assert np.array_equal(ref[0], got[0]) and np.array_equal(ref[1], got[1]) and np.array_equal(ref[2], got[2]) and ref[3] == got[3]


def test_out_of_range_index(emul):
    r, c, v = np.array([0, 5], np.int32), np.array([0, 0], np.int32), np.ones(2)
    assert emul(r, c, v, 5, 5, 0)[3] == -1
