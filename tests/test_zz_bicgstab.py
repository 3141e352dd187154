"""Jacobi-preconditioned BiCGSTAB (the reference's bench_bicg.cpp:150-456): the oracle restatement on the
CPU, and the device-resident implementation (b200spmv_bicgstab) against it on the GPU.  (File name sorts last on purpose:
the newest GPU tests run after the established parity suite.)"""
import numpy as np
import pytest

import matrices as M


def test_oracle_bicgstab_solves_nonsymmetric_system(oracle):
    n = 3000
    rp, col, val = M.dd_matrix(n, 6, 1)
    import scipy.sparse as sp
    A = sp.csr_matrix((val, col, rp), shape=(n, n))
    assert abs(A - A.T).sum() > 1.0                      # really nonsymmetric
    xs = np.random.default_rng(2).uniform(-1, 1, n)
    b = A @ xs
    x, info = oracle.bicgstab(rp, col, val, b, 40)
    assert info["iterations"] == 40                      # no stop test in the reference: the cap IS the iteration count
    assert np.linalg.norm(x - xs) <= 1e-10 * np.linalg.norm(xs)
    assert info["err_best"] <= 1e-12 * np.linalg.norm(b)
    assert info["trace"][0] == pytest.approx(np.linalg.norm(b), rel=1e-14)      # x0 = 0: r0 = b
    # cap 0: x_best = x0 = 0
    x0, i0 = oracle.bicgstab(rp, col, val, b, 0)
    assert i0["iterations"] == 0 and not x0.any()
    # checkpoints: the Laplacian needs ~190 iterations; by 250 the best explicit residual is tiny
    rp, col, val = M.convdiff_matrix(60, 0.0)
    x, info = oracle.bicgstab(rp, col, val, np.ones(3600), 250)
    assert info["err_best"] <= 1e-9 * 60.0
    r = np.ones(3600) - oracle.csr_spmv(rp, col, val, x)
    assert np.linalg.norm(r) == pytest.approx(info["err_best"], rel=1e-6, abs=1e-14)
    # a missing diagonal entry is the reference's "bad K, zero in diagonal"
    rp3, col3, val3 = M.csr_from_lengths([1, 1], 2, 0)
    col3[:] = [1, 0]
    with pytest.raises(ValueError):
        oracle.bicgstab(rp3, col3, val3, np.ones(2), 5)


@pytest.mark.gpu
@pytest.mark.parametrize("fmt_name", ["csr", "sell", "auto"])
def test_device_bicgstab_matches_oracle(native, oracle, fmt_name):
    import spmv_research_b200 as S
    fmt = {"csr": S.CSR_ADAPTIVE, "sell": S.SELL_C_SIGMA, "auto": S.AUTO}[fmt_name]
    for n, per_row, seed in ((3000, 6, 1), (200000, 8, 4)):
        rp, col, val = M.dd_matrix(n, per_row, seed)
        b = np.random.default_rng(seed + 10).uniform(-1, 1, n)
        x_ref, ref = oracle.bicgstab(rp, col, val, b, 40)
        MF = S.csr_to_format(rp, col, val, n, n, int(rp[-1]), format=fmt, keep_diag=1)
        x, res = MF.handle.bicgstab(b, 40, trace=True)
        assert res["iterations"] == 40 and res["products"] == 2 * 40 + 2 and res["restarts"] == 0
        # ||r_k|| of the recurrence: the same sequence while it is above rounding noise (dot products sum in another order)
        tr = res["trace"]
        assert tr.shape == (41,)
        for k in range(12):
            assert abs(tr[k] - ref["trace"][k]) <= 1e-8 * ref["trace"][k], (k, tr[k], ref["trace"][k])
        assert np.linalg.norm(x - x_ref) <= 1e-11 * np.linalg.norm(x_ref)
        # the returned x really solves the system: explicit residual from the CPU oracle's product
        r = b - oracle.csr_spmv(rp, col, val, x)
        assert np.linalg.norm(r) <= 1e-12 * np.linalg.norm(b)
        assert res["err_best"] == pytest.approx(np.linalg.norm(r), rel=1e-3, abs=1e-14 * np.linalg.norm(b))
        # capped run: same iterate as the oracle after 5 steps, within fp64 reordering noise
        x5, r5 = MF.handle.bicgstab(b, 5)
        x5_ref, _ = oracle.bicgstab(rp, col, val, b, 5)
        assert r5["iterations"] == 5 and r5["converged"] == 0
        assert np.linalg.norm(x5 - x5_ref) <= 1e-11 * np.linalg.norm(x5_ref)
        MF.close()


@pytest.mark.gpu
def test_device_bicgstab_checkpoints(native, oracle):
    """More than 100 iterations: the every-100 explicit residual / x_best bookkeeping (bench_bicg.cpp:307-331)."""
    import spmv_research_b200 as S
    nx = 100
    n = nx * nx
    rp, col, val = M.convdiff_matrix(nx, 0.0)
    b = np.ones(n)
    x_ref, ref = oracle.bicgstab(rp, col, val, b, 250)
    MF = S.csr_to_format(rp, col, val, n, n, int(rp[-1]), keep_diag=1)
    x, res = MF.handle.bicgstab(b, 250)
    assert res["iterations"] == 250 and res["products"] == 2 * 250 + 2 + 2
    r = b - oracle.csr_spmv(rp, col, val, x)
    assert np.linalg.norm(r) == pytest.approx(res["err_best"], rel=1e-3, abs=1e-12)
    assert res["err_best"] <= 100.0 * ref["err_best"] + 1e-12 * nx
    assert np.linalg.norm(x - x_ref) <= 1e-8 * np.linalg.norm(x_ref)
    # cap 0 returns x0 = 0 like the reference
    x0, r0 = MF.handle.bicgstab(b, 0)
    assert r0["iterations"] == 0 and not x0.any() and r0["err_best"] == pytest.approx(nx, rel=1e-12)
    MF.close()


@pytest.mark.gpu
def test_device_bicgstab_error_paths(native):
    import spmv_research_b200 as S
    rp, col, val = M.dd_matrix(1000, 4, 9)
    MF = S.csr_to_format(rp, col, val, 1000, 1000, int(rp[-1]))
    with pytest.raises(S.B200Error):
        MF.handle.bicgstab(np.ones(1000), 10)             # created without keep_diag
    MF.close()
    rp3, col3, val3 = M.csr_from_lengths([1, 1], 2, 0)
    col3[:] = [1, 0]
    MF = S.csr_to_format(rp3, col3, val3, 2, 2, 2, keep_diag=1)
    with pytest.raises(S.B200Error, match="zero in diagonal"):
        MF.handle.bicgstab(np.ones(2), 10)
    MF.close()


@pytest.mark.gpu
def test_reference_bicg_harness_drives_the_plugin(native, oracle, tmp_path):
    """The reference's unmodified BiCGSTAB harness (bench.cpp + bench_bicg.cpp) linked with host/b200_plugin.cpp
    (B200SPMV_TRANSFER=always: x changes every call): the ||r_k|| trace of the oracle and a converged x_best."""
    import refharness as RH
    if not RH.available("bicg_b200_d.exe"):
        pytest.skip("oracle/_ref/bicg_b200_d.exe not built (needs /root/reference at build time)")
    n = 3000
    rp, col, val = M.dd_matrix(n, 6, 1)
    path = str(tmp_path / "dd_3000.mtx")
    RH.write_matrix_market(path, rp, col, val, n)
    r = RH.run("bicg_b200_d.exe", [path], gpu_kernel=True, threads=2,
               extra_env={"CG_MAX_NUM_ITERS": "40", "B200SPMV_TRANSFER": "always", "B200SPMV_PIN": "0"})
    assert r["returncode"] == 0, (r["stdout"][-1500:], r["stderr"][-1500:])
    _, info = oracle.bicgstab(rp, col, val, np.ones(n), 40)
    assert len(r["cg_trace"]) == 40
    for k, err, _, _ in r["cg_trace"][:12]:
        assert abs(info["trace"][k] - err) <= 2e-3 * err
    assert r["cg_final_error"] is not None and r["cg_final_error"] <= 1e-12 * np.sqrt(n)
