"""Repro of the slow products seen in the config-3 sweep (profiles/r02b_config3_twins_range.*): the mawi twin's +-30 % variants
-- 14 to 24 million rows of which all but 39 ... 190 000 are empty, five rows of 2 ... 16 million non-zeros -- where CSR took
3 ... 129 ms and format construction 4 s.  Row lengths are the generator's own (pass 1 on the host); columns are synthetic
(the timing question does not depend on them), so the matrix is built in a second instead of the generator's 100 s.

  python tools/repro_giant_rows.py [--sell]           # prints create / product times; run under ncu for the per-kernel list
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import spmv_research_b200 as S  # noqa: E402
from spmv_research_b200 import amg  # noqa: E402

LINES = ["18571154 18571154 2.04836 3932.67 normal random 0.00905 8.00637e+06 0.88395 0.88745 14",
         "18571154 18571154 2.04836 3805.81 normal random 0.00815 8.00637e+06 0.88395 0.88745 14",
         "14237884 14237884 2.04836 3805.81 normal random 0.00905 8.00637e+06 0.88395 0.88745 14"]


def main():
    dev = torch.device("cuda:0")
    fmts = [("csr", S.CSR_ADAPTIVE)] + ([("sell", S.SELL_C_SIGMA)] if "--sell" in sys.argv else [])
    for line in LINES:
        p = amg.MatrixParams.parse(line)
        rp, bw = amg.row_ptr(p)
        m, n, nnz = p.rows, p.cols, int(rp[-1])
        lo = int(n / 2 - bw / 2)
        col = (lo + (np.arange(nnz, dtype=np.int64) * 7919) % max(int(bw), 1)).astype(np.int32)
        val = np.ones(nnz)
        L = np.diff(rp)
        print("== %s\n   nnz %d, non-empty rows %d, rows > 2304: %d, longest %d" % (line, nnz, int((L > 0).sum()), int((L > 2304).sum()), int(L.max())), flush=True)
        dx = torch.ones(n, dtype=torch.float64, device=dev)
        dy = torch.zeros(m, dtype=torch.float64, device=dev)
        for name, fmt in fmts:
            t0 = time.time()
            MF = S.csr_to_format(rp, col, val, m, n, nnz, 0, 0, format=fmt)
            t_create = time.time() - t0
            st = MF.handle.stats()
            ms = MF.handle.time_device(dx, dy, 5, 256 << 20)
            y = dy.cpu().numpy()
            ok = bool(np.array_equal(y, L.astype(np.float64)))          # x = 1, values = 1: y = row length, exact
            print("   %-4s create %.3f s (convert_ms %.1f) product ms %s  launches %d  nsplit %d compressed %d  y == row lengths: %s"
                  % (name, t_create, st["convert_ms"], np.round(ms, 4).tolist(), MF.gpu_launches_per_spmv, int(st["nsplit"]), int(st["compressed_rows"]), ok), flush=True)
            MF.close()


if __name__ == "__main__":
    main()
