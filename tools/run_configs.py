"""BASELINE.json configs 1, 3, 4 and the single-GPU member of 5 on one B200: per matrix, every format's device time
(cold: an L2 flush between launches; warm: back to back), effective GB/s over compulsory bytes, parity of y against
the CPU oracle's serial CSR kernel under the north-star per-row bound (x = random), and the reference's OpenMP CSR
kernel timed on the host cores beside it.

  python tools/run_configs.py [config1] [twins] [g1] [g2] [config5] > gpurun_out/configs.jsonl

One JSON record per matrix on stdout, a readable table on stderr.  Test / measurement tooling: this is one of the
places allowed to load oracle/ (as the checker and the timed CPU baseline), never the product path.
"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
os.environ.setdefault("OMP_DYNAMIC", "false")
import torch  # noqa: E402
import spmv_research_b200 as S  # noqa: E402
from spmv_research_b200 import amg, harness as H  # noqa: E402
from oracle import oracle as O  # noqa: E402

_REAL_STDOUT = os.dup(1)        # the reference's C code printf()s to stdout
os.dup2(2, 1)


def emit(d):
    os.write(_REAL_STDOUT, (json.dumps(d) + "\n").encode())


def workloads(names):
    out = []
    for nm in names:
        if nm == "config1":
            out.append(("config1", H.CONFIG1))
        elif nm == "g1":
            out.append(("config4-G1", H.CONFIG4_G1))
        elif nm == "g2":
            out.append(("config4-G2", H.CONFIG4_G2))
        elif nm == "config5":
            out.append(("config5-1gpu", H.config5(8)))
        elif nm == "twins":
            for p in H.load_param_file(os.path.join(ROOT, "spmv_research_b200", "params", "validation_twins.txt")):
                out.append(("config3-twin:" + p.name, p))
        elif nm == "classes":
            for p in H.load_param_file(os.path.join(ROOT, "spmv_research_b200", "params", "medium_classes.txt")):
                out.append(("config2-class:" + p.name, p))
        elif nm == "medium":
            for p in H.medium_sample():
                out.append(("config2-sample", p))
        else:
            raise SystemExit("unknown workload %s" % nm)
    return out


FORMATS = (("csr", dict(format=S.CSR_ADAPTIVE)), ("sell", dict(format=S.SELL_C_SIGMA)), ("auto", dict(format=S.AUTO)))


def run_one(tag, p, dev, peak, cpu_iters=2):
    t0 = time.time()
    rp, col, val = amg.generate(p)
    t_gen = time.time() - t0
    m, n, nnz = p.rows, p.cols, int(rp[-1])
    bc = H.compulsory_bytes(m, n, nnz)
    lens = np.diff(rp)
    rec = {"config": tag, "matrix": p.line(), "m": m, "n": n, "nnz": nnz, "max_row": int(lens.max()) if m else 0,
           "empty_rows": int((lens == 0).sum()), "compulsory_bytes": int(bc), "generate_s": round(t_gen, 2), "formats": {}}
    x = np.random.default_rng(14).uniform(-1.0, 1.0, n)
    y_ref = O.csr_spmv(rp, col, val, x)
    bound = O.row_abs_bound(rp, col, val, x)
    dx = torch.from_numpy(x).to(dev)
    dy = torch.zeros(m, dtype=torch.float64, device=dev)
    flush = 256 << 20
    for name, kw in FORMATS:
        try:
            MF = S.csr_to_format(rp, col, val, m, n, nnz, 0, 0, **kw)
        except S.B200Error as ex:                       # e.g. SELL padding beyond int32 items on a giant-row twin
            rec["formats"][name] = {"error": str(ex)[:200]}
            continue
        y = np.full(m, 1.0)
        MF.spmv(x, y)
        ratio = float(np.max(np.abs(y - y_ref) / np.maximum(bound, 1e-300))) if m else 0.0
        ok = bool(np.all(np.abs(y - y_ref) <= 1e-12 * bound))
        MF.handle.time_device(dx, dy, 2, flush)
        cold = float(np.median(MF.handle.time_device(dx, dy, 9, flush)))
        warm = float(np.median(MF.handle.time_device(dx, dy, 9, 0)))
        rec["formats"][name] = {"format_name": MF.format_name, "parity_ok": ok, "max_err_over_bound": ratio,
                                "ms_cold": round(cold, 5), "ms_warm": round(warm, 5),
                                "gflops_cold": round(2.0 * nnz / cold * 1e-6, 1), "eff_GBs_cold": round(bc / cold * 1e-6, 1),
                                "frac_of_measured_peak_cold": round(bc / cold * 1e-6 / peak, 4),
                                "gflops_warm": round(2.0 * nnz / warm * 1e-6, 1), "launches": MF.gpu_launches_per_spmv,
                                "convert_ms": round(MF.convert_ms, 2), "trial_ms": MF.trial_ms}
        MF.close()
    # the reference's OpenMP CSR kernel on the host cores (reported baseline)
    if O.ref_available():
        R = O.RefCSR(rp, col, val, n)
        yc = np.full(m, 1.0)
        R.spmv(x, yc)
        t0 = time.perf_counter()
        for _ in range(cpu_iters):
            R.spmv(x, yc)
        tc = (time.perf_counter() - t0) / cpu_iters
        rec["cpu_reference"] = {"gflops": round(2.0 * nnz / tc * 1e-9, 3), "ms": round(tc * 1e3, 3), "cores": O.max_threads(),
                                "bit_equal_to_serial_oracle": bool(np.array_equal(yc, y_ref))}
    del dx, dy
    torch.cuda.empty_cache()
    return rec


def main():
    names = sys.argv[1:] or ["config1", "twins", "g1", "g2"]
    dev = torch.device("cuda:0")
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6650.0
    print("%-34s %10s %11s %8s | %9s %9s %9s | %6s %6s | %8s %s" % ("config", "rows", "nnz", "maxrow", "csr ms", "sell ms", "auto ms",
                                                                    "frac", "GF/s", "cpu GF/s", "parity"), file=sys.stderr)
    for tag, p in workloads(names):
        rec = run_one(tag, p, dev, peak)
        emit(rec)
        f = rec["formats"]
        g = lambda k, q: f.get(k, {}).get(q, float("nan"))
        print("%-34s %10d %11d %8d | %9.4f %9.4f %9.4f | %6.3f %6.0f | %8.2f %s" % (
            tag[:34], rec["m"], rec["nnz"], rec["max_row"], g("csr", "ms_cold"), g("sell", "ms_cold"), g("auto", "ms_cold"),
            g("auto", "frac_of_measured_peak_cold"), g("auto", "gflops_cold"), rec.get("cpu_reference", {}).get("gflops", float("nan")),
            "ok" if all(v.get("parity_ok", False) for v in f.values() if "error" not in v) else "FAIL"), file=sys.stderr, flush=True)


if __name__ == "__main__":
    main()
