"""fp32 (`_f` build axis) beside fp64 on the GPU box: for config 1 and the medium sample, the fp32 CSR product
(b200spmv_f32_*) is checked against the pinned `_f` oracle under the fp32 per-row bound, then timed on the device (CUDA events
on the launching stream inside the library, cold: an L2 flush before every product) next to the fp64 CSR kernel on the same
matrix.  One JSON line per matrix + a summary line.  roofline figures use each precision's own compulsory bytes:
  B_c(f64) = 12 nnz + 12 m + 8 n + 4,   B_c(f32) = 8 nnz + 8 m + 4 n + 4.
  python tools/run_f32.py [--small]"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

FLUSH = 256 << 20          # > 126 MB L2


def peak_gbs():
    try:
        return float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        return 6549.8       # B200_PROFILING.md fallback (measured copy bandwidth)


def main():
    import torch
    import spmv_research_b200 as S
    from spmv_research_b200 import amg, harness
    from oracle import oracle as O          # checker only
    small = "--small" in sys.argv
    dev = "cuda:0"
    lines = ["200000 200000 10 5 normal random 0.3 0 0.5 0.5 14"] if small else \
        ["1000000 1000000 10 5 normal random 0.3 0 0.5 0.5 14"] + [p.line() for p in harness.medium_sample()]
    peak = peak_gbs()
    iters = 5 if small else 20
    tot = {"f32_ms": 0.0, "f64_ms": 0.0, "f32_bc": 0, "f64_bc": 0, "nnz": 0}
    for line in lines:
        p = amg.MatrixParams.parse(line)
        rp, col, val = amg.generate(p)
        m, n, nnz = p.rows, p.cols, int(rp[-1])
        x32 = np.random.default_rng(14).uniform(-1.0, 1.0, n).astype(np.float32)
        MF = S.csr_to_format_f32(rp, col, val, m, n, nnz)
        y = np.full(m, 1.0, np.float32)
        MF.spmv(x32, y)
        y_ref = O.csr_spmv_f32(rp, col, val, x32)
        bound = O.row_abs_bound_f32(rp, col, val, x32)
        tol = 2.0 * (np.diff(rp) + 2) * 2.0 ** -24 * bound           # tests/test_f32_emul.py::f32_tolerance
        err = np.abs(y.astype(np.float64) - y_ref.astype(np.float64))
        ok = bool(np.all(err <= tol))
        dx32 = torch.from_numpy(x32).to(dev)
        dy32 = torch.empty(m, dtype=torch.float32, device=dev)
        MF.handle.time_device(dx32, dy32, 3, FLUSH)                  # warm-up
        ms32 = float(np.median(MF.handle.time_device(dx32, dy32, iters, FLUSH)))
        bc32 = MF.compulsory_bytes()
        MF.close()
        M64 = S.csr_to_format(rp, col, val, m, n, nnz, 0, 0, format=S.CSR_ADAPTIVE)
        dx64 = torch.from_numpy(x32.astype(np.float64)).to(dev)
        dy64 = torch.empty(m, dtype=torch.float64, device=dev)
        M64.handle.time_device(dx64, dy64, 3, FLUSH)
        ms64 = float(np.median(M64.handle.time_device(dx64, dy64, iters, FLUSH)))
        bc64 = harness.compulsory_bytes(m, n, nnz)
        M64.close()
        for k, v in (("f32_ms", ms32), ("f64_ms", ms64), ("f32_bc", bc32), ("f64_bc", bc64), ("nnz", nnz)):
            tot[k] += v
        print(json.dumps({"matrix": line, "nnz": nnz, "parity_f32": ok, "worst_err_over_tol": float(np.max(err / np.maximum(tol, 1e-300))),
                          "f32": {"ms": round(ms32, 5), "gflops": round(2e-6 * nnz / ms32, 1), "eff_GBs": round(bc32 / ms32 * 1e-6, 1),
                                  "frac_of_measured": round(bc32 / ms32 * 1e-6 / peak, 4)},
                          "f64": {"ms": round(ms64, 5), "gflops": round(2e-6 * nnz / ms64, 1), "eff_GBs": round(bc64 / ms64 * 1e-6, 1),
                                  "frac_of_measured": round(bc64 / ms64 * 1e-6 / peak, 4)},
                          "speedup_f32_over_f64": round(ms64 / ms32, 3)}), flush=True)
        del dx32, dy32, dx64, dy64
        torch.cuda.empty_cache()
    print(json.dumps({"summary": True, "matrices": len(lines), "timing": "median of %d cold products (256 MB L2 flush before each), CUDA events" % iters,
                      "peak_GBs": peak,
                      "f32": {"gflops": round(2e-6 * tot["nnz"] / tot["f32_ms"], 1), "eff_GBs": round(tot["f32_bc"] / tot["f32_ms"] * 1e-6, 1),
                              "frac_of_measured": round(tot["f32_bc"] / tot["f32_ms"] * 1e-6 / peak, 4)},
                      "f64": {"gflops": round(2e-6 * tot["nnz"] / tot["f64_ms"], 1), "eff_GBs": round(tot["f64_bc"] / tot["f64_ms"] * 1e-6, 1),
                              "frac_of_measured": round(tot["f64_bc"] / tot["f64_ms"] * 1e-6 / peak, 4)}}), flush=True)


if __name__ == "__main__":
    main()
