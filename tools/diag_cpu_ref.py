"""Diagnostic: the reference OpenMP CSR kernel timed in one process before / after importing torch and after CUDA
initialisation, with y verified against the serial oracle each time (is a fast number a complete product?)."""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from spmv_research_b200 import amg, harness as H  # noqa: E402
from oracle import oracle as O  # noqa: E402

print("omp max threads", O.max_threads(), "cpu_count", os.cpu_count(), "affinity", len(os.sched_getaffinity(0)), "OMP env",
      {k: v for k, v in os.environ.items() if k.startswith(("OMP", "GOMP", "KMP", "MKL"))})
mats = []
for p in H.medium_sample()[:3]:
    rp, col, val = amg.generate(p)
    mats.append((p, rp, col, val, O.RefCSR(rp, col, val, p.cols), O.csr_spmv(rp, col, val, np.ones(p.cols))))


def leg(tag):
    for p, rp, col, val, R, yser in mats:
        x = np.ones(p.cols)
        y = np.full(p.rows, 1.0)
        R.spmv(x, y)
        t0 = time.perf_counter()
        for _ in range(3):
            R.spmv(x, y)
        t = (time.perf_counter() - t0) / 3
        print("%-22s %-60s %8.3f ms %7.2f GFLOP/s  y == serial oracle: %s  affinity %d" % (
            tag, p.line()[:60], t * 1e3, 2 * int(rp[-1]) / t * 1e-9, bool(np.array_equal(y, yser)), len(os.sched_getaffinity(0))), flush=True)


leg("no torch")
import torch  # noqa: E402
leg("torch imported")
torch.zeros(1, device="cuda")
torch.cuda.synchronize()
leg("CUDA initialised")
import spmv_research_b200 as S  # noqa: E402
p, rp, col, val, R, yser = mats[0]
MF = S.csr_to_format(rp, col, val, p.rows, p.cols, int(rp[-1]))
leg("after csr_to_format")
