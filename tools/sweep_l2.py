"""x band larger than L2: sweep of the L2 budget (fraction of the band gathered with evict_last) on one B200.
  python tools/sweep_l2.py > gpurun_out/sweep_l2.txt"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import spmv_research_b200 as S  # noqa: E402
from spmv_research_b200 import amg, harness as H  # noqa: E402

LINES = ["4194304 33554432 60 20 normal random 0.3 0 0.5 0.5 14",            # a config-5 shard: band 80 MB
         "55042369 55042369 2.1295885720 0.6731069663 normal random 0.3370633097 17.3133965463 0.0573248727 0.1163647359 14",   # kmer twin: 148 MB
         "28508159 28508159 5 1.6667 normal random 0.6 0 0.5 0.5 14",        # largest medium class: 137 MB
         "13418495 13418495 10 3.3333 normal random 0.6 0 0.5 0.5 14",       # 64 MB
         "4329604 4329604 10 3.3333 normal random 0.6 100 0.95 0.5 14"]      # 21 MB: control (fraction 1 at every budget >= 21)
dev = torch.device("cuda:0")
print("%-100s %-5s %7s %6s %9s %8s" % ("matrix", "fmt", "x_l2_mb", "frac", "ms", "GB/s"))
for ln in LINES:
    p = amg.MatrixParams.parse(ln)
    rp, col, val = amg.generate(p)
    m, n, nnz = p.rows, p.cols, int(rp[-1])
    bc = H.compulsory_bytes(m, n, nnz)
    dx = torch.ones(n, dtype=torch.float64, device=dev)
    dy = torch.zeros(m, dtype=torch.float64, device=dev)
    for fmt, fname in ((S.CSR_ADAPTIVE, "csr"), (S.SELL_C_SIGMA, "sell")):
        for mb in ((56, 1), (56, 2), (56, 3), (56, 4), (56, 0)) if os.environ.get("SWEEP_PANELS") else [(v, 1) for v in (4000, 112, 84, 64, 56, 48, 36, 24)]:
            mb, panels = mb
            MF = S.csr_to_format(rp, col, val, m, n, nnz, 0, 0, format=fmt, col_panels=panels)
            lo, hi = MF.handle.col_range()
            frac = min(1.0, mb * 2 ** 20 / (8.0 * (hi - lo)))
            MF.handle.time_device(dx, dy, 3, 256 << 20)
            ms = float(np.median(MF.handle.time_device(dx, dy, 9, 256 << 20)))
            print("%-100s %-5s %7d %6.3f %9.4f %8.0f  panels req %d -> %d" % (ln[:100], fname, mb, frac, ms, bc / ms * 1e-6, panels, MF.col_panels), flush=True)
            MF.close()
    del dx, dy
    torch.cuda.empty_cache()
