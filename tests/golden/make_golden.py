"""Regenerates tests/golden/ref_golden.npz from the REFERENCE's own code (oracle/_ref/*.so built by
oracle/Makefile from /root/reference).  Run in the authoring container only:

    python tests/golden/make_golden.py

The fixture pins (a) the generator (FNV-1a-64 of row_ptr / col_ind / values), (b) the nnz-balanced
partitioner, (c) merge-path coordinates, (d) the SELL-C-sigma layout, (e) y of the serial CSR kernel.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from oracle import oracle as O  # noqa: E402
import matrices as M  # noqa: E402
from spmv_research_b200.amg import MatrixParams  # noqa: E402

PART_W = [1, 2, 3, 4, 8, 61, 1024]
MERGE_T = [1, 3, 8, 64]
SELL_CFG = [(32, 16384), (32, 64), (4, 8), (256, 16384), (8, 8), (2, 4096)]


def main():
    assert O.ref_available(), "build oracle/_ref first (make -C oracle)"
    out = {}
    for li, line in enumerate(M.GOLDEN_LINES):
        p = MatrixParams.parse(line)
        rp, col, val = O.ref_amg_generate(p.rows, p.cols, p.avg, p.std, p.distribution, p.placement, p.bw, p.skew, p.neigh, p.crs, p.seed)
        m, nnz = p.rows, int(rp[-1])
        k = "g%d_" % li
        out[k + "nnz"] = np.int64(nnz)
        out[k + "fnv"] = np.array([O.fnv1a64(rp), O.fnv1a64(col), O.fnv1a64(val)], np.uint64)
        for W in PART_W:
            out[k + "part%d" % W] = O.ref_partition(W, rp, m, nnz)
        for T in MERGE_T:
            per = (m + nnz + T - 1) // T
            out[k + "merge%d" % T] = np.array([O.ref_merge_path_search(min(per * t, m + nnz), rp, m, nnz) for t in range(T + 1)], np.int64)
        for (C_, sg) in SELL_CFG:
            S = O.ref_sellcs_convert(rp, col, val, p.cols, C_, sg, index64=True)
            out[k + "sell_%d_%d" % (C_, sg)] = np.array([O.fnv1a64(S["row_order"]), O.fnv1a64(S["slice_width"]), O.fnv1a64(S["slice_ptr"]),
                                                         O.fnv1a64(S["col"]), O.fnv1a64(S["val"])], np.uint64)
        x = M.rand_x(p.cols)
        y = np.full(m, 1.0)
        O.RefCSR(rp, col, val, p.cols).spmv_serial(x, y)
        # keep the fixture small: every 7th row of y plus a checksum of all of it
        out[k + "y_sample"] = y[::7].copy()
        out[k + "y_fnv"] = np.uint64(O.fnv1a64(y))
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "ref_golden.npz")
    np.savez_compressed(path, **out)
    print("wrote", path, os.path.getsize(path), "bytes,", len(out), "arrays")


if __name__ == "__main__":
    main()
