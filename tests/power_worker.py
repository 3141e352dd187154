"""One rank of a multi-GPU run of the C power layer (b200spmv_power_*), started by tests/test_zz_power_c.py (and by hand:
`python tests/power_worker.py --rank R --world W --uid-file F --out OUT.npz ...` once per rank).

Per-step parity as SURVEY.md 8e asks: before each checked iteration the GPU's x_k and scale are copied to the host, the
oracle's serial CSR kernel (subkernel_csr_scalar, K/csr.cpp:334-350) is applied to this rank's shard, and the GPU's own slot
of x_{k+1} must lie within the per-row bound |y - y_ref| <= 1e-12 * sum_j |a_ij x_j|; the band rows of x_{k+1} as held by
THIS rank are written out so that the parent can require them bit-equal on every rank (what the peers received)."""
import argparse
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--rank", type=int, required=True)
    ap.add_argument("--world", type=int, required=True)
    ap.add_argument("--uid-file", required=True)
    ap.add_argument("--out", required=True)
    ap.add_argument("--line", default="400000 400000 30 10 normal random 0.3 0 0.5 0.5 14")
    ap.add_argument("--exchange", default="fused")
    ap.add_argument("--scalar-sync", type=int, default=0)
    ap.add_argument("--chunks", type=int, default=0)
    ap.add_argument("--panels", type=int, default=1)
    ap.add_argument("--format", default="csr")
    ap.add_argument("--graph", type=int, default=0)
    ap.add_argument("--steps", type=int, default=4)
    args = ap.parse_args()
    import spmv_research_b200 as S
    from spmv_research_b200 import amg, capi, power
    from oracle import oracle as O          # checker
    rank, world = args.rank, args.world
    capi.check(capi.lib().b200spmv_set_device(rank))
    p = amg.MatrixParams.parse(args.line)
    rp, bw = amg.row_ptr(p)
    m, n, nnz = p.rows, p.cols, int(rp[-1])
    bounds = power.shard_bounds(rp, m, nnz, world, capi.partition)
    i_s, i_e = int(bounds[rank, 0]), int(bounds[rank, 1])
    col, val = amg.fill(p, rp, bw, i_s, i_e)
    rp_loc = (rp[i_s:i_e + 1] - rp[i_s]).astype(np.int32)
    fmt = {"csr": S.CSR_ADAPTIVE, "sell": S.SELL_C_SIGMA}[args.format]
    MF = S.csr_to_format(rp_loc, col, val, i_e - i_s, n, int(rp_loc[-1]), 0, 0, format=fmt, device=rank, col_panels=args.panels)
    uid = None
    if world > 1:
        if rank == 0:
            uid = capi.comm_unique_id()
            with open(args.uid_file + ".tmp", "wb") as f:
                f.write(uid)
            os.rename(args.uid_file + ".tmp", args.uid_file)
        else:
            t0 = time.time()
            while not os.path.exists(args.uid_file):
                if time.time() - t0 > 120:
                    raise SystemExit("rank %d: no NCCL id after 120 s" % rank)
                time.sleep(0.05)
            uid = open(args.uid_file, "rb").read()
    it = power.CPowerIteration(MF, bounds, n, rank, world, rank, exchange=args.exchange, chunks=args.chunks, use_graph=bool(args.graph),
                               scalar_sync=args.scalar_sync, uid=uid)
    info = it.info()
    x0 = np.random.default_rng(5).uniform(0.5, 1.5, n)
    it.set_start(x0)
    worst, bands, eigs = 0.0, [], []
    lo, hi = info["band_lo"], info["band_hi"]
    for k in range(args.steps):
        xk, s = it.get_x()
        if k == 0:
            assert np.array_equal(xk, x0)
        # columns outside this shard's referenced range are never read: neutralise whatever stale value they hold
        c_lo, c_hi = MF.handle.col_range()
        xr = np.zeros(n)
        xr[c_lo:c_hi] = xk[c_lo:c_hi]
        y_ref = s * O.csr_spmv(rp_loc, col, val, xr)
        bound = abs(s) * O.row_abs_bound(rp_loc, col, val, xr)
        it.iterate(1)
        xn, s2 = it.get_x()
        err = np.abs(xn[i_s:i_e] - y_ref)
        worst = max(worst, float(np.max(err / (1e-12 * np.maximum(bound, 1e-300)))))
        nrm = float(np.sqrt(np.sum(xn[i_s:i_e] ** 2)))
        bands.append(xn[lo:hi].copy())
        eigs.append(it.eigenvalue())
        np.testing.assert_allclose(1.0 / s2, it.eigenvalue(), rtol=1e-13)
        if world == 1:
            np.testing.assert_allclose(nrm, it.eigenvalue(), rtol=1e-12)
    # a longer run through the graph path (if asked), then one more checked step
    if args.graph:
        it.iterate(10)
        xk, s = it.get_x()
        c_lo, c_hi = MF.handle.col_range()
        xr = np.zeros(n)
        xr[c_lo:c_hi] = xk[c_lo:c_hi]
        y_ref = s * O.csr_spmv(rp_loc, col, val, xr)
        bound = abs(s) * O.row_abs_bound(rp_loc, col, val, xr)
        it.iterate(1)
        xn, _ = it.get_x()
        worst = max(worst, float(np.max(np.abs(xn[i_s:i_e] - y_ref) / (1e-12 * np.maximum(bound, 1e-300)))))
        bands.append(xn[lo:hi].copy())
        eigs.append(it.eigenvalue())
        info = it.info()
    np.savez(args.out, worst=worst, bands=np.stack(bands), eigs=np.array(eigs), info=np.array([info[k] for k in sorted(info)]),
             info_keys=np.array(sorted(info)), slot=np.array([i_s, i_e]))
    it.close()
    MF.close()
    print("rank %d ok: worst err / (1e-12 bound) = %.3g, eig = %.15g, info = %s" % (rank, worst, eigs[-1], info), flush=True)


if __name__ == "__main__":
    main()
