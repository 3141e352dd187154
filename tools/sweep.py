"""Variant sweep on one B200: every kernel option x a few generator matrices, CUDA-event timing
through b200spmv_time_device (an L2 flush between launches), effective GB/s over compulsory bytes.
  python tools/sweep.py [quick|full] > gpurun_out/sweep.txt
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import spmv_research_b200 as S  # noqa: E402
from spmv_research_b200 import amg, capi, harness as H  # noqa: E402
import torch  # noqa: E402

mode = sys.argv[1] if len(sys.argv) > 1 else "quick"
lines = [H.CONFIG1.line()] + [p.line() for p in H.medium_sample()]
if mode == "full":
    lines += ["8388607 8388607 50 16.6667 gamma random 0.6 10000 0.05 0.05 14", "8388607 8388607 50 150 gamma random 0.6 0 0.05 0.05 14",
              "4194304 4194304 60 20 normal random 0.3 0 0.5 0.5 14"]
variants = [("csr npt8", dict(format=S.CSR_ADAPTIVE, csr_nnz_per_thread=8)),
            ("sell s16384 v0", dict(format=S.SELL_C_SIGMA, sell_sigma=16384)),
            ("sell s16384 v1", dict(format=S.SELL_C_SIGMA, sell_sigma=16384, sell_variant=1)),
            ("sell s16384 v2", dict(format=S.SELL_C_SIGMA, sell_sigma=16384, sell_variant=2)),
            ("sell s16384 v3", dict(format=S.SELL_C_SIGMA, sell_sigma=16384, sell_variant=3)),
            ("sell s256 v0", dict(format=S.SELL_C_SIGMA, sell_sigma=256)),
            ("sell s256 v1", dict(format=S.SELL_C_SIGMA, sell_sigma=256, sell_variant=1)),
            ("sell s1024 v1", dict(format=S.SELL_C_SIGMA, sell_sigma=1024, sell_variant=1)),
            ("sell s64 v1", dict(format=S.SELL_C_SIGMA, sell_sigma=64, sell_variant=1)),
            ("auto", dict(format=S.AUTO))]
if os.environ.get("SWEEP_CSR"):
    variants = [("csr npt4", dict(format=S.CSR_ADAPTIVE, csr_nnz_per_thread=4)), ("csr npt8", dict(format=S.CSR_ADAPTIVE, csr_nnz_per_thread=8)),
                ("csr npt16", dict(format=S.CSR_ADAPTIVE, csr_nnz_per_thread=16)), ("csr npt8 nohint", dict(format=S.CSR_ADAPTIVE, csr_nnz_per_thread=8, x_hint=1))]
dev = torch.device("cuda:0")
print("%-72s %-18s %9s %9s %8s %8s %9s" % ("matrix", "variant", "ms(med)", "ms(min)", "GB/s", "GFLOP/s", "conv ms"))
for ln in lines:
    p = amg.MatrixParams.parse(ln)
    rp, col, val = amg.generate(p)
    m, n, nnz = p.rows, p.cols, int(rp[-1])
    bc = H.compulsory_bytes(m, n, nnz)
    dx = torch.ones(n, dtype=torch.float64, device=dev)
    dy = torch.zeros(m, dtype=torch.float64, device=dev)
    flush = 256 << 20 if bc < (1 << 30) else 0          # matrices below 1 GB: rewrite 256 MB between launches
    for name, kw in variants:
        MF = S.csr_to_format(rp, col, val, m, n, nnz, 0, 0, **kw)
        MF.handle.time_device(dx, dy, 3, flush)
        ms = MF.handle.time_device(dx, dy, 15, flush)
        med, mn = float(np.median(ms)), float(ms.min())
        print("%-72s %-18s %9.4f %9.4f %8.0f %8.0f %9.2f %s" % (ln, name, med, mn, bc / med * 1e-6, 2.0 * nnz / med * 1e-6, MF.convert_ms,
                                                              (MF.format_name + " " + str(MF.trial_ms)) if name == "auto" else ""), flush=True)
        MF.close()
    del dx, dy
    torch.cuda.empty_cache()
