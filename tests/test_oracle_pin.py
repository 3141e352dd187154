"""Pins the CPU oracle (oracle/spmv_oracle.c): against the committed golden fixture made from the
reference's own code (tests/golden/make_golden.py) and, when oracle/_ref is present, against that
code directly on random cases with empty rows, duplicate prefix sums and ragged tails."""
import os

import numpy as np
import pytest

import matrices as M
from tests.golden.make_golden import MERGE_T, PART_W, SELL_CFG


def test_partition_vs_golden(oracle, golden, golden_matrices):
    for li, (p, rp, col, val) in enumerate(golden_matrices):
        nnz = int(rp[-1])
        for W in PART_W:
            assert np.array_equal(oracle.partition(W, rp, p.rows, nnz), golden["g%d_part%d" % (li, W)]), (li, W)


def test_merge_path_vs_golden(oracle, golden, golden_matrices):
    for li, (p, rp, col, val) in enumerate(golden_matrices):
        m, nnz = p.rows, int(rp[-1])
        for T in MERGE_T:
            per = (m + nnz + T - 1) // T
            got = np.array([oracle.merge_path_search(min(per * t, m + nnz), rp, m, nnz) for t in range(T + 1)], np.int64)
            assert np.array_equal(got, golden["g%d_merge%d" % (li, T)]), (li, T)
            _, coords = oracle.merge_csrmv(T, rp, col, val, M.rand_x(p.cols))
            assert np.array_equal(coords, got)


def test_sellcs_vs_golden(oracle, golden, golden_matrices):
    for li, (p, rp, col, val) in enumerate(golden_matrices):
        for (C_, sg) in SELL_CFG:
            S = oracle.sellcs_convert(rp, col, val, C_, sg)
            got = np.array([oracle.fnv1a64(S[k]) for k in ("row_order", "slice_width", "slice_ptr", "col", "val")], np.uint64)
            assert np.array_equal(got, golden["g%d_sell_%d_%d" % (li, C_, sg)]), (li, C_, sg)


def test_csr_scalar_vs_golden(oracle, golden, golden_matrices):
    """The serial kernel is pinned up to FMA contraction of the reference build: within the
    north-star bound, and bit-equal wherever the compiler made the same choice."""
    for li, (p, rp, col, val) in enumerate(golden_matrices):
        x = M.rand_x(p.cols)
        y = oracle.csr_spmv(rp, col, val, x)
        bound = oracle.row_abs_bound(rp, col, val, x)
        assert np.all(np.abs(y[::7] - golden["g%d_y_sample" % li]) <= 1e-12 * bound[::7])


def test_sell_and_merge_spmv_agree_with_csr(oracle, golden_matrices):
    for (p, rp, col, val) in golden_matrices[:4]:
        x = M.rand_x(p.cols)
        y = oracle.csr_spmv(rp, col, val, x)
        bound = oracle.row_abs_bound(rp, col, val, x)
        ys = oracle.sellcs_spmv(oracle.sellcs_convert(rp, col, val, 32, 16384), x)
        ym, _ = oracle.merge_csrmv(7, rp, col, val, x)
        assert oracle.within_bound(ys, y, bound) and oracle.within_bound(ym, y, bound)


def test_gold_f128_and_rowblocks(oracle):
    name, rp, col, val, n = M.edge_cases()[4]
    x = M.rand_x(n)
    y = oracle.csr_spmv(rp, col, val, x)
    acc = oracle.check_accuracy(rp, col, val, x, y)
    assert acc["max_rel"] < 1e-10 and acc["max_ae"] < 1e-12          # bench_spmv.cpp:114-119 pass criterion
    rb = oracle.adaptive_rowblocks(rp, 64, 4)
    assert rb[0] == 0 and rb[-1] == len(rp) - 1 and np.all(np.diff(rb) >= 0) and (len(rb) - 1) % 4 == 0
    for a, b in zip(rb[:-1], rb[1:]):
        assert b - a <= 1 or rp[b] - rp[a] <= 64                   # multi-row blocks hold <= BLOCK_SIZE nnz


# ---- direct comparison with the compiled reference (authoring container; skipped when _ref is absent)
def _need_ref(oracle):
    if not oracle.ref_available():
        pytest.skip("oracle/_ref not built")


def test_partition_vs_reference_random(oracle):
    _need_ref(oracle)
    rng = np.random.default_rng(3)
    for _ in range(300):
        m = int(rng.integers(1, 400))
        lens = rng.integers(0, 9, m) * (rng.random(m) < rng.random())
        rp = np.zeros(m + 1, np.int32)
        np.cumsum(lens, out=rp[1:])
        rp += int(rng.integers(0, 5))                      # Sums may start anywhere (parallel_util.h:150)
        total = int(rp[-1] - rp[0])
        W = int(rng.integers(1, 40))
        assert np.array_equal(oracle.partition(W, rp, m, total), oracle.ref_partition(W, rp, m, total))


def test_merge_search_vs_reference_random(oracle):
    _need_ref(oracle)
    rng = np.random.default_rng(4)
    for _ in range(60):
        m = int(rng.integers(1, 300))
        rp, col, val = M.csr_from_lengths(rng.integers(0, 7, m) * (rng.random(m) < 0.7), 50, int(rng.integers(1 << 30)))
        nnz = int(rp[-1])
        for d in rng.integers(0, m + nnz + 1, 25):
            assert oracle.merge_path_search(int(d), rp, m, nnz) == oracle.ref_merge_path_search(int(d), rp, m, nnz)
        x = M.rand_x(50)
        T = int(rng.integers(1, 9))
        y, _ = oracle.merge_csrmv(T, rp, col, val, x)
        yr = oracle.ref_merge_csrmv(T, rp, col, val, x)
        assert oracle.within_bound(y, yr, oracle.row_abs_bound(rp, col, val, x))


def test_sellcs_vs_reference_random(oracle):
    _need_ref(oracle)
    rng = np.random.default_rng(5)
    for it in range(40):
        m = int(rng.integers(1, 700))
        C_ = int(2 ** rng.integers(1, 6))
        sg = C_ * int(rng.integers(1, 9))
        rp, col, val = M.csr_from_lengths(rng.integers(0, 20, m) * (rng.random(m) < 0.75), 90, it)
        A = oracle.sellcs_convert(rp, col, val, C_, sg)
        for i64 in (True, False):
            B = oracle.ref_sellcs_convert(rp, col, val, 90, C_, sg, index64=i64)
            for k in ("row_order", "slice_width", "slice_ptr", "col", "val"):
                assert np.array_equal(A[k], B[k]), (it, k, i64)


def test_csr_vs_reference_plugin(oracle):
    _need_ref(oracle)
    for name, rp, col, val, n in M.edge_cases():
        if len(rp) - 1 < 1 or rp[-1] == 0:
            continue
        x = M.rand_x(n)
        R = oracle.RefCSR(rp, col, val, n)
        assert R.format_name == "Custom_CSR_B"
        yr = np.full(len(rp) - 1, 1.0)
        R.spmv_serial(x, yr)
        yo = oracle.csr_spmv(rp, col, val, x)
        assert oracle.within_bound(yo, yr, oracle.row_abs_bound(rp, col, val, x)), name
        yp = np.full(len(rp) - 1, 1.0)
        R.spmv(x, yp)                                      # OpenMP compute_csr, nnz-balanced threads
        assert np.array_equal(yp, yr), name


def test_pcg_vs_reference_cg_harness(oracle, tmp_path):
    """oracle_pcg pinned against the reference's OWN CG harness (bench.cpp + bench_cg.cpp + K/csr.cpp, compiled unmodified
    into oracle/_ref/cg_csr_d.exe): same eps, same iteration count, the same ||r_k|| at every iteration (the harness prints
    4 significant digits; its dot products are OpenMP-reduced, so later iterates agree to rounding noise only) and the same
    explicit residual at exit.  b = ones as in bench_cg.cpp:512-516 when no _b.mtx file exists."""
    import refharness as RH
    if not RH.available("cg_csr_d.exe"):
        pytest.skip("oracle/_ref/cg_csr_d.exe not built (needs /root/reference at build time)")
    for n, per_row, seed, cap in ((3000, 6, 1, 500), (800, 10, 2, 500), (3000, 6, 1, 7)):
        rp, col, val = M.spd_matrix(n, per_row, seed)
        path = str(tmp_path / ("spd_%d_%d.mtx" % (n, seed)))
        RH.write_matrix_market(path, rp, col, val, n)
        r = RH.run("cg_csr_d.exe", [path], gpu_kernel=False, threads=2, extra_env={"CG_MAX_NUM_ITERS": str(cap)})
        assert r["returncode"] == 0, r["stderr"][-1500:]
        trace = r["cg_trace"]
        x, info = oracle.pcg(rp, col, val, np.ones(n), cap)
        # the harness prints one line per iteration it EXECUTES (k = 0 .. iterations - 1)
        assert len(trace) == info["iterations"], (len(trace), info)
        # replay the oracle with growing caps to read its ||r_k|| sequence: `err` is evaluated at the top of an iteration
        # (bench_cg.cpp:213-218), so a run capped at k + 1 iterations leaves err = ||r_k||, the value the harness printed at k
        for k, err, _, _ in trace[:12]:
            if k + 1 > info["iterations"]:
                break
            _, ik = oracle.pcg(rp, col, val, np.ones(n), k + 1)
            assert ik["iterations"] == k + 1
            assert abs(ik["err"] - err) <= 2e-3 * err + 1e-300, (k, ik["err"], err)
        assert abs(info["err_explicit"] - r["cg_final_error"]) <= 0.05 * r["cg_final_error"] + 1e-13 * np.sqrt(n)


def test_bicgstab_vs_reference_bicg_harness(oracle, tmp_path):
    """oracle_bicgstab pinned against the reference's OWN BiCGSTAB harness (bench.cpp + bench_bicg.cpp + K/csr.cpp, compiled
    unmodified into oracle/_ref/bicg_csr_d.exe): it runs exactly CG_MAX_NUM_ITERS iterations (no stop test), prints ||r_k|| at
    the top of every iteration (4 significant digits) and the explicit residual of the returned x_best at exit."""
    import refharness as RH
    if not RH.available("bicg_csr_d.exe"):
        pytest.skip("oracle/_ref/bicg_csr_d.exe not built (needs /root/reference at build time)")
    cases = (("dd", M.dd_matrix(3000, 6, 1), 3000, 40), ("lap", M.convdiff_matrix(60, 0.0), 3600, 230),
             ("cd", M.convdiff_matrix(60, 0.05), 3600, 7))
    for tag, (rp, col, val), n, cap in cases:
        path = str(tmp_path / ("bicg_%s.mtx" % tag))
        RH.write_matrix_market(path, rp, col, val, n)
        r = RH.run("bicg_csr_d.exe", [path], gpu_kernel=False, threads=2, extra_env={"CG_MAX_NUM_ITERS": str(cap)})
        assert r["returncode"] == 0, r["stderr"][-1500:]
        trace = r["cg_trace"]
        x, info = oracle.bicgstab(rp, col, val, np.ones(n), cap)
        assert len(trace) == cap == info["iterations"]
        for k, err, _, _ in trace[:25]:
            assert abs(info["trace"][k] - err) <= 2e-3 * err + 1e-300, (tag, k, info["trace"][k], err)
        # checkpoints (k = 100, 200): the harness prints error_explicit / error_best from then on
        if cap > 200:
            assert trace[-1][3] < trace[0][3]
            assert abs(np.log10(trace[-1][3] / info["err_best"])) < 1.0 or info["err_best"] <= trace[-1][3]
        # the harness's closing "error =" is ||b - A x_best||; the oracle's err_best is the same quantity
        assert abs(info["err_best"] - r["cg_final_error"]) <= 0.25 * r["cg_final_error"] + 1e-13 * np.sqrt(n), (tag, info, r["cg_final_error"])


def test_plain_sell_vs_reference(oracle):
    """the oracle's restatement of the reference's plain SELL (K/sell.cpp) pinned against that file compiled
    unmodified (oracle/_ref/libref_sell.so): row_cluster_ptr / ja / a bit for bit -- including the padding entries, which carry
    the last real column the converting THREAD saw and therefore change with the thread count -- and the fused-multiply-add
    product bit for bit, tail lanes of the last slice included."""
    import os
    if not os.path.exists(os.path.join(os.path.dirname(oracle.__file__), "_ref", "libref_sell.so")):
        pytest.skip("oracle/_ref/libref_sell.so not built (needs /root/reference at build time)")
    rng = np.random.default_rng(5)
    cases = [M.csr_from_lengths(rng.integers(0, 9, 1003), 500, 1),                          # ragged last slice, empty rows
             M.csr_from_lengths(np.r_[np.zeros(9, np.int64), rng.integers(0, 40, 500)], 2000, 2),   # a thread range starting on empty rows
             M.csr_from_lengths(np.r_[rng.integers(1, 4, 300), [700], rng.integers(1, 4, 200)], 900, 3),   # one long row
             M.csr_from_lengths([3], 10, 4)]                                                   # a single row
    t0 = oracle.max_threads()
    try:
        for nt in (1, 2, 3, 5):
            oracle.set_threads(nt)
            for rp, col, val in cases:
                m, n = len(rp) - 1, int(col.max()) + 1 if len(col) else 1
                R = oracle.RefSELL(rp, col, val, n)
                O = oracle.sell_plain_convert(rp, col, val, R.VL, nt)
                assert np.array_equal(O["row_cluster_ptr"], R.row_cluster_ptr)
                assert np.array_equal(O["ja"], R.ja), (nt, m)
                assert np.array_equal(O["a"], R.a)
                x = M.rand_x(n, 7)
                y_ref = np.full(((m + R.VL - 1) // R.VL) * R.VL + 64, 3.0)
                R.spmv(x, y_ref)
                y = oracle.sell_plain_spmv(O, x)
                assert np.array_equal(y, y_ref[:len(y)]) and np.all(y_ref[len(y):] == 3.0)
                # against the scalar CSR oracle: same sums up to the fused multiply-add (north-star bound)
                assert oracle.within_bound(y[:m], oracle.csr_spmv(rp, col, val, x), oracle.row_abs_bound(rp, col, val, x))
    finally:
        oracle.set_threads(t0)


def test_sorted_sell_vs_reference(oracle):
    """SURVEY 8a row a8, second half: the per-thread-sorted SELL (K/sell_sorted.cpp) restated and pinned against that file
    compiled unmodified (oracle/_ref/libref_sell_sorted.so): permutation / rev_permutation (stable ascending sort inside each
    converting thread's row range), row_cluster_ptr / ja / a bit for bit, and the product (y written through rev_permutation)
    bit-equal to the plain SELL oracle's.  Sizes keep every thread at >= VEC_LEN slices: with fewer the reference's
    `ii_e_last -= VEC_LEN` (:356-358) walks out of its range."""
    import os
    if not os.path.exists(os.path.join(os.path.dirname(oracle.__file__), "_ref", "libref_sell_sorted.so")):
        pytest.skip("oracle/_ref/libref_sell_sorted.so not built (needs /root/reference at build time)")
    rng = np.random.default_rng(6)
    cases = [M.csr_from_lengths(rng.integers(0, 9, 1003), 500, 1),
             M.csr_from_lengths(np.r_[np.zeros(9, np.int64), rng.integers(0, 40, 500)], 2000, 2),
             M.csr_from_lengths(np.r_[rng.integers(1, 4, 300), [700], rng.integers(1, 4, 201)], 900, 3)]
    t0 = oracle.max_threads()
    try:
        for nt in (1, 2, 3, 5):
            oracle.set_threads(nt)
            for rp, col, val in cases:
                m, n = len(rp) - 1, int(col.max()) + 1
                R = oracle.RefSELLSorted(rp, col, val, n)
                O = oracle.sell_sorted_convert(rp, col, val, R.VL, nt)
                assert np.array_equal(O["permutation"], R.permutation), (nt, m)
                assert np.array_equal(O["rev_permutation"], R.rev_permutation)
                assert np.array_equal(O["row_cluster_ptr"], R.row_cluster_ptr)
                assert np.array_equal(O["ja"], R.ja), (nt, m)
                assert np.array_equal(O["a"], R.a)
                x = M.rand_x(n, 7)
                y_ref = np.full(m + 64, 3.0)
                R.spmv(x, y_ref)
                y = oracle.sell_sorted_spmv(O, x)
                assert np.array_equal(y, y_ref[:m]) and np.all(y_ref[m:] == 3.0)
                P = oracle.sell_plain_convert(rp, col, val, R.VL, nt)
                assert np.array_equal(y, oracle.sell_plain_spmv(P, x)[:m])
    finally:
        oracle.set_threads(t0)


def test_fp32_serial_csr_vs_reference_f_build(oracle):
    """oracle_csr_spmv_scalar_f32 against the reference's `_f` build of K/csr.cpp (ValueType = float,
    make.sh:215-216) compiled unmodified into oracle/_ref/libref_csr_f.so -- the serial kernel bit for bit (float products,
    one float accumulator, ascending j, double values narrowed by the constructor's cast), and the OpenMP product, which is
    the same kernel per thread range."""
    if not oracle.ref_available() or not os.path.exists(os.path.join(os.path.dirname(oracle.__file__), "_ref", "libref_csr_f.so")):
        pytest.skip("oracle/_ref/libref_csr_f.so not built")
    rng = np.random.default_rng(11)
    for m, n, per in [(1, 1, 1), (300, 500, 7), (5000, 4000, 40), (64, 100000, 3000)]:
        lens = rng.integers(0, 2 * per + 1, m)
        rp, col, val = M.csr_from_lengths(lens, n, 5)
        val = val * 10.0 ** rng.integers(-6, 6, val.size)           # narrowing to float must round, not just truncate zeros
        x = rng.uniform(-1, 1, n).astype(np.float32)
        R = oracle.RefCSRf32(rp, col, val, n)
        y_ser = np.full(m, 1.0, np.float32)
        R.spmv_serial(x, y_ser)
        y_omp = np.full(m, 1.0, np.float32)
        R.spmv(x, y_omp)
        y = oracle.csr_spmv_f32(rp, col, val, x)
        assert y.dtype == np.float32
        assert np.array_equal(y, y_ser) and np.array_equal(y, y_omp), (m, n, per)
        b = oracle.row_abs_bound_f32(rp, col, val, x)
        assert np.all(np.abs(y.astype(np.float64) - oracle.csr_spmv(rp, col, val.astype(np.float32).astype(np.float64), x.astype(np.float64)))
                      <= (lens + 2) * 2.0 ** -24 * b + 1e-300)


def test_reference_f_harness_error_level(oracle, tmp_path):
    """The reference's harness built on its `_f` axis (DOUBLE=0, ValueType=float; oracle/_ref/spmv_csr_f.exe) with its own CPU
    CSR plugin: runs, reports float-level errors against the _Float128 gold -- and trips its own 1e-7 relative criterion
    (bench_spmv.cpp:116,198).  The fp32 oracle agrees with what it computes (same mae to the printed digits), which is the
    level tests/test_zz_f32_device.py holds the GPU plugin to."""
    from spmv_research_b200 import amg
    import refharness as RH
    if not RH.available("spmv_csr_f.exe"):
        pytest.skip("oracle/_ref/spmv_csr_f.exe not built (needs /root/reference at build time)")
    p = amg.MatrixParams.parse("30000 30000 12 6 normal random 0.3 50 0.5 0.5 14")
    rp, col, val = amg.generate(p)
    path = str(tmp_path / "amg_30000.mtx")
    RH.write_matrix_market(path, rp, col, val, p.cols)
    r = RH.run("spmv_csr_f.exe", [path], gpu_kernel=False, threads=2)
    assert r["returncode"] == 0, r["stderr"][-2000:]
    assert any("Custom_CSR_B" in f for f in r["csv"]), r["csv"]
    assert 1e-9 < r["errors"]["mae"] < 1e-5                      # float, not double (the `_d` build gives ~1e-16)
    # the harness multiplies by x = ones (bench_spmv.cpp:598-606): the oracle's `_f` kernel gives the same mean absolute error
    y32 = oracle.csr_spmv_f32(rp, col, val, np.ones(p.cols, np.float32)).astype(np.float64)
    gold = np.array([np.sum(np.sort(val[rp[i]:rp[i + 1]].astype(np.float32).astype(np.float64))) for i in range(p.rows)])
    mae = np.mean(np.abs(y32 - gold))
    assert abs(mae - r["errors"]["mae"]) <= 0.02 * r["errors"]["mae"], (mae, r["errors"]["mae"])
