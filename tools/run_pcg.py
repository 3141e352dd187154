"""Device-resident Jacobi-PCG (b200spmv_pcg) next to the CPU oracle's restatement of the reference's preconditioned_cg
on one SPD system: iterations, time per iteration, agreement of the solutions.  python tools/run_pcg.py [rows] [per_row]"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import matrices as M  # noqa: E402
import spmv_research_b200 as S  # noqa: E402
from oracle import oracle as O  # noqa: E402

n = int(sys.argv[1]) if len(sys.argv) > 1 else 2000000
per_row = int(sys.argv[2]) if len(sys.argv) > 2 else 8
rp, col, val = M.spd_matrix(n, per_row, 5, dominance=1.0005)
nnz = int(rp[-1])
b = np.random.default_rng(6).uniform(-1, 1, n)
t0 = time.time()
x_ref, ref = O.pcg(rp, col, val, b, 2000)
t_cpu = time.time() - t0
print("system: %d rows, %d nnz; oracle (serial CPU): %d iterations in %.2f s (%.2f ms / iteration)" % (n, nnz, ref["iterations"], t_cpu, 1e3 * t_cpu / max(ref["iterations"], 1)))
for name, fmt in (("csr", S.CSR_ADAPTIVE), ("sell", S.SELL_C_SIGMA), ("auto", S.AUTO)):
    MF = S.csr_to_format(rp, col, val, n, n, nnz, format=fmt, keep_diag=1)
    MF.handle.pcg(b, 5)
    t0 = time.time()
    x, res = MF.handle.pcg(b, 2000)
    t = time.time() - t0
    print("%-5s %-28s %d iterations (%d products, %d restarts) in %.4f s = %.4f ms / iteration incl. H2D b, D2H x; converged %d; "
          "|x - x_oracle| / |x| = %.2e; explicit residual / |b| = %.2e" % (
              name, MF.format_name, res["iterations"], res["products"], res["restarts"], t, 1e3 * t / max(res["iterations"], 1), res["converged"],
              np.linalg.norm(x - x_ref) / np.linalg.norm(x_ref), res["err_best"] / np.linalg.norm(b)), flush=True)
    MF.close()
