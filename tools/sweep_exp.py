"""Experimental builds of the CSR kernel's row reduction against the shipped library, on one B200.

The reduction stage costs the CSR kernel ~30 % over its stream + gather floor while a conflict-free shared-memory round
trip costs 14 % (profiles/r01a_kexp_floors.txt): the difference is mostly bank conflicts of the thread-per-row
reads.  spmv_research_b200/build.py::EXPERIMENTS compiles the same library with csr_kernel.cu's switches (LDS.128 pair
reads; at least 2 / 4 / 8 lanes per row instead of thread per row; the conflict-free row -> thread table of csr_perm.h, whose
y must be bit-identical to the shipped library's, as must the pair reads'); this tool runs each library in its own process
(B200SPMV_LIB), checks y against the oracle's serial CSR kernel under the per-row bound, and times cold products.

  python tools/sweep_exp.py                 # parent: builds are expected in lib/exp/ (python -m ... build_experiments here)
  python tools/sweep_exp.py --child TAG     # one library: JSON lines
"""
import json
import os
import subprocess
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)

LINES = ["1000000 1000000 10 5 normal random 0.3 0 0.5 0.5 14"]       # + the medium sample, added in child()


def child(tag):
    import torch
    import spmv_research_b200 as S
    from spmv_research_b200 import amg, harness as H
    from oracle import oracle as O          # checker only
    dev = torch.device("cuda:0")
    flush = 256 << 20
    for ln in LINES + [p.line() for p in H.medium_sample()]:
        p = amg.MatrixParams.parse(ln)
        rp, col, val = amg.generate(p)
        m, n, nnz = p.rows, p.cols, int(rp[-1])
        MF = S.csr_to_format(rp, col, val, m, n, nnz, 0, 0, format=S.CSR_ADAPTIVE)
        x = np.random.default_rng(14).uniform(-1.0, 1.0, n)
        y = np.full(m, 1.0)
        MF.spmv(x, y)
        y_ref = O.csr_spmv(rp, col, val, x)
        ok = bool(np.all(np.abs(y - y_ref) <= 1e-12 * O.row_abs_bound(rp, col, val, x)))
        dx = torch.ones(n, dtype=torch.float64, device=dev)
        dy = torch.zeros(m, dtype=torch.float64, device=dev)
        MF.handle.time_device(dx, dy, 3, flush)
        ms = MF.handle.time_device(dx, dy, 25, flush)
        bc = H.compulsory_bytes(m, n, nnz)
        print(json.dumps({"lib": tag, "matrix": ln, "parity": ok, "y_fnv": int(O.fnv1a64(y)), "ms_med": round(float(np.median(ms)), 5), "ms_min": round(float(ms.min()), 5),
                          "eff_GBs": round(bc / float(np.median(ms)) * 1e-6, 1)}), flush=True)
        MF.close()
        del dx, dy


def main():
    if len(sys.argv) > 2 and sys.argv[1] == "--child":
        return child(sys.argv[2])
    from spmv_research_b200 import build as B
    libs = {"shipped": None}
    exp_dir = os.path.join(ROOT, "spmv_research_b200", "lib", "exp")
    for tag in sorted(B.EXPERIMENTS):
        so = os.path.join(exp_dir, "libb200spmv_%s.so" % tag)
        if os.path.exists(so):
            libs[tag] = so
    res = {}
    order = list(libs.items()) + [("shipped", None)]      # the shipped library first and last, so that drift of the box shows
    for rep, (tag, so) in enumerate(order):
        if True:
            env = dict(os.environ)
            if so:
                env["B200SPMV_LIB"] = so
            else:
                env.pop("B200SPMV_LIB", None)
            out = subprocess.run([sys.executable, os.path.abspath(__file__), "--child", tag], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE,
                                 text=True, timeout=900)
            if out.returncode != 0:
                print(json.dumps({"lib": tag, "error": out.stderr[-400:]}), flush=True)
                continue
            for ln in out.stdout.splitlines():
                r = json.loads(ln)
                r["run"] = rep
                print(json.dumps(r), flush=True)
                res.setdefault(tag, {}).setdefault(r["matrix"], []).append(r)
    base = res.get("shipped", {})
    if base and all(len(rs) > 1 for rs in base.values()):
        print(json.dumps({"drift_shipped_last_over_first": round(sum(rs[-1]["ms_med"] for rs in base.values()) / sum(rs[0]["ms_med"] for rs in base.values()), 4)}),
              flush=True)
    for tag, mats in res.items():
        tot = sum(min(r["ms_med"] for r in rs) for rs in mats.values())
        tot0 = sum(min(r["ms_med"] for r in base[k]) for k in mats if k in base)
        print(json.dumps({"summary": tag, "matrices": len(mats), "parity_all": all(r["parity"] for rs in mats.values() for r in rs),
                          "y_bit_identical_to_shipped": all(r["y_fnv"] == base[k][0]["y_fnv"] for k, rs in mats.items() if k in base for r in rs),
                          "sum_ms": round(tot, 4), "vs_shipped": round(tot / tot0, 4) if tot0 else None,
                          "per_matrix_vs_shipped": {k: round(min(r["ms_med"] for r in rs) / min(r["ms_med"] for r in base[k]), 3)
                                                    for k, rs in mats.items() if k in base}}), flush=True)


if __name__ == "__main__":
    main()
