"""The edge-case set of tests/matrices.py through every product path, small enough to run under compute-sanitizer
(tools/gpu_sanitize.sh: memcheck and racecheck): CSR (plain, compressed rows), SELL-C-sigma (several C / sigma), column
panels, the fused epilogue (scale, norm, peer destinations -- a second local buffer stands in for the peer mapping), the
host-vector path with its chunked D2H overlap, fp32, COO -> CSR, the device generator, PCG.  Prints one line per case;
y is checked against the oracle (checker only) so that a sanitizer-clean but wrong run is not mistaken for a pass."""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch  # noqa: E402
import matrices as M  # noqa: E402
import spmv_research_b200 as S  # noqa: E402
from spmv_research_b200 import amg, capi, formats  # noqa: E402
from oracle import oracle as O  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    st = torch.cuda.current_stream().cuda_stream
    bad = 0
    cases = [(name, rp, col, val, n) for name, rp, col, val, n in M.edge_cases()]
    for line in M.GOLDEN_LINES[:4]:
        p = amg.MatrixParams.parse(line)
        rp, col, val = amg.generate(p)
        cases.append((line.split()[0] + "x" + line.split()[2][:4], rp, col, val, p.cols))
    for name, rp, col, val, n in cases:
        m, nnz = len(rp) - 1, int(rp[-1])
        x = M.rand_x(n, 3)
        y_ref = O.csr_spmv(rp, col, val, x)
        bound = O.row_abs_bound(rp, col, val, x)
        variants = [("csr", dict(format=S.CSR_ADAPTIVE)), ("csr_compressed", dict(format=S.CSR_ADAPTIVE, csr_compress_rows=2)),
                    ("sell", dict(format=S.SELL_C_SIGMA)), ("sell_c8_s64", dict(format=S.SELL_C_SIGMA, sell_C=8, sell_sigma=64)),
                    ("panels3", dict(format=S.CSR_ADAPTIVE, col_panels=3)), ("panels2_sell", dict(format=S.SELL_C_SIGMA, col_panels=2))]
        for vname, kw in variants:
            MF = S.csr_to_format(rp, col, val, m, n, nnz, 0, 0, **kw)
            y = np.full(m + 8, 1.0)
            MF.spmv(x, y)
            ok = bool(np.all(np.abs(y[:m] - y_ref) <= 1e-12 * bound)) and bool(np.all(y[m:] == 1.0))
            # fused epilogue: scale, norm, one "peer" destination limited to a row range
            dx = torch.from_numpy(x).to(dev)
            dy = torch.full((max(m, 1),), 1.0, dtype=torch.float64, device=dev)
            peer = torch.zeros(max(m, 1), dtype=torch.float64, device=dev)
            scale = torch.tensor([0.5], dtype=torch.float64, device=dev)
            norm2 = torch.zeros(1, dtype=torch.float64, device=dev)
            MF.spmv_device(dx, dy, st, scale=scale, norm2=norm2, peers=[peer], peer_rows=(m // 4, m // 2 + 1))
            torch.cuda.synchronize()
            yd = dy.cpu().numpy()[:m]
            ok = ok and bool(np.all(np.abs(yd - 0.5 * y_ref) <= 1e-12 * bound))
            MF.close()
            bad += 0 if ok else 1
            print("%-32s %-16s m=%-6d nnz=%-8d %s" % (name, vname, m, nnz, "ok" if ok else "MISMATCH"), flush=True)
        if nnz > 0:
            F = formats.csr_to_format_f32(rp, col, val, m, n, nnz)
            yf = np.full(m, 1.0, np.float32)
            F.spmv(x.astype(np.float32), yf)
            F.close()
            # COO -> CSR on the device from a shuffled copy
            perm = np.random.default_rng(1).permutation(nnz)
            rows = np.repeat(np.arange(m, dtype=np.int32), np.diff(rp))[perm]
            MFc = formats.coo_to_format(rows, col[perm], val[perm], m, n)
            y = np.full(m, 1.0)
            MFc.spmv(x, y)
            ok = bool(np.all(np.abs(y - y_ref) <= 1e-12 * bound))
            MFc.close()
            bad += 0 if ok else 1
            print("%-32s %-16s %s" % (name, "f32 + coo", "ok" if ok else "MISMATCH"), flush=True)
    # device generator (warp-per-segment kernel) on two small lines
    for line in ("20000 20000 10 5 normal random 0.3 0 0.5 0.5 14", "6000 6000 50 16.6667 gamma random 0.6 1000 0.05 0.05 14"):
        p = amg.MatrixParams.parse(line)
        rp, col, val = amg.generate(p)
        _, d_rp, d_col, d_val = amg.generate_device(p)
        ok = bool(np.array_equal(d_col.cpu().numpy(), col)) and bool(np.array_equal(d_val.cpu().numpy(), val))
        bad += 0 if ok else 1
        print("%-32s %-16s %s" % (line.split()[0], "device generator", "ok" if ok else "MISMATCH"), flush=True)
    # PCG (dot / update kernels) on a small SPD matrix
    rps, cols, vals = M.spd_matrix(3000, 6, 8)
    b = np.random.default_rng(9).uniform(-1, 1, 3000)
    MF = S.csr_to_format(rps, cols, vals, 3000, 3000, int(rps[-1]), keep_diag=1)
    xs, res = MF.handle.pcg(b, 200)
    MF.close()
    print("pcg converged=%d iterations=%d" % (res["converged"], res["iterations"]), flush=True)
    bad += 0 if res["converged"] == 1 else 1
    print("cases with a wrong result: %d" % bad)
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
