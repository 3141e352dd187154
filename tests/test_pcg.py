"""Jacobi-preconditioned CG: the oracle restatement of bench_cg.cpp:93-320 on the CPU, and the
device-resident implementation (b200spmv_pcg) against it on the GPU."""
import numpy as np
import pytest

import matrices as M


def test_oracle_pcg_solves_spd_system(oracle):
    rp, col, val = M.spd_matrix(3000, 6, 1)
    import scipy.sparse as sp
    A = sp.csr_matrix((val, col, rp), shape=(3000, 3000))
    xs = np.random.default_rng(2).uniform(-1, 1, 3000)
    b = A @ xs
    x, info = oracle.pcg(rp, col, val, b, 500)
    assert 0 < info["iterations"] < 500 and info["err"] < info["eps"]
    assert np.linalg.norm(x - xs) <= 1e-10 * np.linalg.norm(xs)
    assert info["err_best"] <= 1e-12 * np.linalg.norm(b)
    # iteration cap: returns x_best of the last explicit residual, err above eps
    x2, info2 = oracle.pcg(rp, col, val, b, 3)
    assert info2["iterations"] == 3 and info2["err"] > info2["eps"]
    # a missing diagonal entry is the reference's "bad K, zero in diagonal"
    rp3, col3, val3 = M.csr_from_lengths([1, 1], 2, 0)
    col3[:] = [1, 0]
    with pytest.raises(ValueError):
        oracle.pcg(rp3, col3, val3, np.ones(2), 5)


@pytest.mark.gpu
@pytest.mark.parametrize("fmt_name", ["csr", "sell", "auto"])
def test_device_pcg_matches_oracle(native, oracle, fmt_name):
    import spmv_research_b200 as S
    fmt = {"csr": S.CSR_ADAPTIVE, "sell": S.SELL_C_SIGMA, "auto": S.AUTO}[fmt_name]
    for n, per_row, seed in ((5000, 5, 3), (200000, 8, 4)):
        rp, col, val = M.spd_matrix(n, per_row, seed)
        b = np.random.default_rng(seed + 10).uniform(-1, 1, n)
        x_ref, ref = oracle.pcg(rp, col, val, b, 1000)
        MF = S.csr_to_format(rp, col, val, n, n, int(rp[-1]), format=fmt, keep_diag=1)
        x, res = MF.handle.pcg(b, 1000)
        assert res["converged"] == 1 and res["err"] < res["eps"]
        assert abs(res["iterations"] - ref["iterations"]) <= 3, (res, ref)          # summation order differs in the dot products
        assert res["products"] == res["iterations"] + 2 + (res["iterations"] - 1) // 100
        assert np.linalg.norm(x - x_ref) <= 1e-11 * np.linalg.norm(x_ref)
        # the returned x really solves the system: explicit residual from the CPU oracle's product
        r = b - oracle.csr_spmv(rp, col, val, x)
        assert np.linalg.norm(r) <= 1e-12 * np.linalg.norm(b)
        # capped run: same iterate as the oracle after 7 steps, within fp64 reordering noise
        x7, r7 = MF.handle.pcg(b, 7)
        x7_ref, _ = oracle.pcg(rp, col, val, b, 7)
        assert r7["iterations"] == 7 and r7["converged"] == 0
        assert np.linalg.norm(x7 - x7_ref) <= 1e-11 * np.linalg.norm(x7_ref)
        MF.close()


@pytest.mark.gpu
def test_device_pcg_error_paths(native):
    import spmv_research_b200 as S
    rp, col, val = M.spd_matrix(1000, 4, 9)
    MF = S.csr_to_format(rp, col, val, 1000, 1000, int(rp[-1]))
    with pytest.raises(S.B200Error):
        MF.handle.pcg(np.ones(1000), 10)                  # created without keep_diag
    MF.close()
    rp3, col3, val3 = M.csr_from_lengths([1, 1], 2, 0)
    col3[:] = [1, 0]
    MF = S.csr_to_format(rp3, col3, val3, 2, 2, 2, keep_diag=1)
    with pytest.raises(S.B200Error, match="zero in diagonal"):
        MF.handle.pcg(np.ones(2), 10)
    MF.close()


@pytest.mark.gpu
def test_reference_cg_harness_drives_the_plugin(native, oracle, tmp_path):
    """The reference's unmodified CG harness (bench.cpp + bench_cg.cpp) linked with host/b200_plugin.cpp: its
    preconditioned_cg() calls MF->spmv with a different x every iteration, so the plugin runs with B200SPMV_TRANSFER=always
    (the reference's own GPU plugins cache x and fail here).  Same iteration count as the oracle (+- 2), converged."""
    import refharness as RH
    if not RH.available("cg_b200_d.exe"):
        pytest.skip("oracle/_ref/cg_b200_d.exe not built (needs /root/reference at build time)")
    n = 3000
    rp, col, val = M.spd_matrix(n, 6, 1)
    path = str(tmp_path / "spd_3000.mtx")
    RH.write_matrix_market(path, rp, col, val, n)
    r = RH.run("cg_b200_d.exe", [path], gpu_kernel=True, threads=2,
               extra_env={"CG_MAX_NUM_ITERS": "500", "B200SPMV_TRANSFER": "always", "B200SPMV_PIN": "0"})
    assert r["returncode"] == 0, (r["stdout"][-1500:], r["stderr"][-1500:])
    _, info = oracle.pcg(rp, col, val, np.ones(n), 500)
    assert abs(len(r["cg_trace"]) - info["iterations"]) <= 2, (len(r["cg_trace"]), info)
    assert r["cg_final_error"] is not None and r["cg_final_error"] <= 1e-12 * np.sqrt(n)
