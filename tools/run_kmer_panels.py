"""the kmer_V2a twin (55 M rows, 2.1 non-zeros per row, x band 148 MB > L2) ran at 0.196 of the measured HBM
peak with 4.9 x the compulsory DRAM traffic.  Column panels localise the gathers at the price of one read-modify-write of y per
extra panel; with row compression allowed inside panels (round 2) a sparse panel only touches the rows it has entries in.
This sweeps the panel count x row compression on that matrix (generated on the GPU), cold device times, y checked against the
single-panel product under the per-row bound (x = 1, values >= 0: the bound is y itself).

  python tools/run_kmer_panels.py > gpurun_out/kmer_panels.jsonl
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
from spmv_research_b200 import amg, capi, formats, harness as H  # noqa: E402

LINE = "55042369 55042369 2.1295885720 0.6731069663 normal random 0.3370633097 17.3133965463 0.0573248727 0.1163647359 14"


def main():
    dev = torch.device("cuda:0")
    p = amg.MatrixParams.parse(sys.argv[1] if len(sys.argv) > 1 else LINE)
    rp, d_rp, d_col, d_val = amg.generate_device(p)
    m, n, nnz = p.rows, p.cols, int(rp[-1])
    bc = H.compulsory_bytes(m, n, nnz)
    try:
        peak = float(json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))["hbm_gbs"])
    except Exception:
        peak = 6650.0
    dx = torch.ones(n, dtype=torch.float64, device=dev)
    y_ref = None
    for fmt_name, fmt in (("csr", capi.CSR_ADAPTIVE), ("sell", capi.SELL_C_SIGMA)):
        for panels in (1, 2, 3, 4, 6, 8):
            for comp in ((0, 2) if fmt == capi.CSR_ADAPTIVE else (0,)):
                if panels == 1 and comp == 2:
                    continue
                try:
                    MF = formats.csr_to_format_device(d_rp, d_col, d_val, m, n, nnz, format=fmt, col_panels=panels, csr_compress_rows=comp)
                except capi.B200Error as ex:
                    print(json.dumps({"format": fmt_name, "panels": panels, "compress": comp, "error": str(ex)[:200]}), flush=True)
                    continue
                dy = torch.full((m,), 1.0, dtype=torch.float64, device=dev)
                MF.handle.time_device(dx, dy, 2, 256 << 20)
                ms = float(np.median(MF.handle.time_device(dx, dy, 7, 256 << 20)))
                if y_ref is None:
                    y_ref = dy.clone()
                ok = bool(torch.all((dy - y_ref).abs() <= 2e-12 * y_ref.abs()).item())
                print(json.dumps({"format": fmt_name, "panels": panels, "compress": {0: "auto", 2: "always"}[comp], "ms_cold": round(ms, 4),
                                  "eff_GBs": round(bc / ms * 1e-6, 1), "frac_of_measured_peak": round(bc / ms * 1e-6 / peak, 4),
                                  "convert_ms": round(MF.convert_ms, 1), "parity_vs_single_panel": ok}), flush=True)
                MF.close()
                del dy
    print(json.dumps({"matrix": p.line(), "nnz": nnz, "compulsory_bytes": int(bc), "x_band_MB": round(8 * p.bw * n / 2 ** 20, 1)}))


if __name__ == "__main__":
    main()
