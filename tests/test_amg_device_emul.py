"""Device pass 2 of the generator (csrc/amg_fill.h), checked WITHOUT a GPU: the very source the kernel is
built from is compiled as plain C++ with a one-lane warp (tests/emul/amg_fill_emul.cpp -- test infrastructure) and must
reproduce the host generator (csrc/amg.c, itself pinned on the reference generator) bit for bit: whole matrices, arbitrary row
ranges (segments sticking out of the range go through scratch), every placement, giant skew rows (global-memory sort), rows
around the staging / insertion-sort boundaries."""
import ctypes as C
import os
import subprocess

import numpy as np
import pytest

import matrices as M
from spmv_research_b200 import amg

HERE = os.path.dirname(os.path.abspath(__file__))
SRC = os.path.join(HERE, "emul", "amg_fill_emul.cpp")
SO = os.path.join(HERE, "emul", "libamg_fill_emul.so")
CSRC = os.path.join(HERE, "..", "spmv_research_b200", "csrc")

LINES = list(M.GOLDEN_LINES) + [
    "1000 1000 5 1.6667 normal random 0.001 0 0.05 0.5 14",
    "5000 5000 20 6 normal simple 0.05 0 1.9 0.95 3",
    "2500 4000 8 30 gamma diagonal 0.9 3 0.0 0.0 99",
    "700 700 3 0 normal random 2.0 0 0 0 1",
    "60000 60000 50 16.6667 gamma random 0.6 10000 0.05 0.05 14",      # one 60 000-long row: sorted in "global memory"
    "6000 6000 500 166 normal random 0.05 1000 1.9 0.95 14",           # rows around AMGD_STAGE = 512
    "8000 8000 12 2 normal diagonal 0.3 0 0.95 0.95 7",                # rows around AMGD_SMALL = 12
]


@pytest.fixture(scope="module")
def emul(native):
    deps = [SRC, os.path.join(CSRC, "amg_fill.h"), os.path.join(CSRC, "amg_fill_plan.h")]
    if not os.path.exists(SO) or any(os.path.getmtime(d) > os.path.getmtime(SO) for d in deps):
        subprocess.check_call(["g++", "-O2", "-ffp-contract=off", "-fPIC", "-shared", "-Wall", "-Wextra", "-o", SO, SRC])
    L = C.CDLL(SO)
    L.amgd_emul_fill.restype = C.c_int64
    L.amgd_emul_fill.argtypes = [C.c_int64, C.c_int64, C.c_char, C.c_double, C.c_double, C.c_uint32, C.c_void_p, C.c_double,
                                 C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]
    L.amgd_emul_sort.argtypes = [C.c_void_p, C.c_int64]

    def fill(p, rp, bw, a, b):
        cnt = int(rp[b]) - int(rp[a])
        col = np.full(cnt, -7, np.int32)
        val = np.full(cnt, -7.0)
        rc = L.amgd_emul_fill(p.rows, p.cols, p.placement[:1].encode(), p.neigh, p.crs, p.seed & 0xFFFFFFFF, rp.ctypes.data, bw,
                              a, b, col.ctypes.data, val.ctypes.data)
        assert rc == 0, rc
        return col, val
    fill.lib = L
    return fill


@pytest.mark.parametrize("line", LINES)
def test_device_source_reproduces_host_generator(emul, line):
    p = amg.MatrixParams.parse(line)
    rp, bw = amg.row_ptr(p)
    col, val = amg.fill(p, rp, bw, 0, p.rows)
    c2, v2 = emul(p, rp, bw, 0, p.rows)
    assert np.array_equal(col, c2) and np.array_equal(val, v2)
    cuts = [0, 1, p.rows // 3 + 5, p.rows // 2, p.rows - 1, p.rows]
    for a, b in zip(cuts[:-1], cuts[1:]):
        c, v = emul(p, rp, bw, a, b)
        assert np.array_equal(c, col[rp[a]:rp[b]]) and np.array_equal(v, val[rp[a]:rp[b]]), (a, b)


def test_device_source_vs_reference_generator(emul, oracle):
    if not oracle.ref_available():
        pytest.skip("oracle/_ref not built")
    for line in LINES[-7:-3]:
        p = amg.MatrixParams.parse(line)
        rp_ref, col_ref, val_ref = oracle.ref_amg_generate(p.rows, p.cols, p.avg, p.std, p.distribution, p.placement, p.bw, p.skew,
                                                           p.neigh, p.crs, p.seed)
        rp, bw = amg.row_ptr(p)
        assert np.array_equal(rp, rp_ref)
        c, v = emul(p, rp, bw, 0, p.rows)
        assert np.array_equal(c, col_ref) and np.array_equal(v, val_ref), line


def test_warp_sort_network(emul):
    """The same-direction bitonic network with virtual +inf padding sorts every length (not only powers of two)."""
    rng = np.random.default_rng(0)
    for n in list(range(0, 70)) + [127, 128, 129, 511, 512, 513, 1000, 4097, 100000]:
        a = rng.permutation(max(n, 1) * 3)[:n].astype(np.int32)
        b = a.copy()
        emul.lib.amgd_emul_sort(b.ctypes.data, n)
        assert np.array_equal(b, np.sort(a)), n


def test_plan_bounds():
    """Period and set capacity follow the reference's anchor rule (artificial_matrix_generation.c:259-262)."""
    for m, want in [(100, 100), (1000, 100), (17099, 100), (1000000, 244), (33554432, 8192)]:
        anchors = min(m // 100, 4096)
        assert m // anchors == want
