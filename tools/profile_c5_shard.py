"""One rank's shard of BASELINE config 5 at N = 8 (4.19 M of 33.5 M rows, 252 M non-zeros, x band 80 MB) on one GPU, for ncu:
the plain product with and without column panels.  Round 1's capture of the unpanelled shard read 5.48 GB from DRAM for 3.34 GB
of compulsory bytes (L2 hit 34 %, profiles/r01o_ncu_full_c5shard_csr_key_metrics.txt); this is the same shard with 2 panels.

  ncu --set full --clock-control none -k regex:csr_tile_kernel -s 4 -c 2 -o prof python tools/profile_c5_shard.py --panels 2
"""
import argparse
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import spmv_research_b200 as S  # noqa: E402
from spmv_research_b200 import amg, capi, formats, harness as H, power  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--panels", type=int, default=2)
    ap.add_argument("--world", type=int, default=8)
    ap.add_argument("--rank", type=int, default=0)
    ap.add_argument("--iters", type=int, default=5)
    args = ap.parse_args()
    dev = torch.device("cuda:0")
    p = H.config5(args.world)
    rp, bw = amg.row_ptr(p)
    m, n, nnz = p.rows, p.cols, int(rp[-1])
    bounds = power.shard_bounds(rp, m, nnz, args.world, capi.partition)
    i_s, i_e = int(bounds[args.rank, 0]), int(bounds[args.rank, 1])
    rp_loc = (rp[i_s:i_e + 1] - rp[i_s]).astype(np.int32)
    nnz_loc = int(rp_loc[-1])
    d_rp = torch.from_numpy(rp).to(dev)
    col = torch.empty(nnz_loc, dtype=torch.int32, device=dev)
    val = torch.empty(nnz_loc, dtype=torch.float64, device=dev)
    amg.fill_device(p, d_rp, bw, i_s, i_e, col, val, torch.cuda.current_stream().cuda_stream)
    MF = formats.csr_to_format_device(torch.from_numpy(rp_loc).to(dev), col, val, i_e - i_s, n, nnz_loc, format=S.CSR_ADAPTIVE, col_panels=args.panels)
    dx = torch.ones(n, dtype=torch.float64, device=dev)
    dy = torch.zeros(i_e - i_s, dtype=torch.float64, device=dev)
    ms = MF.handle.time_device(dx, dy, args.iters, 0)
    bc = H.compulsory_bytes(i_e - i_s, n, nnz_loc)
    print("shard rows %d nnz %d compulsory bytes %d panels %d: ms per product %s (back to back, x band %.0f MB)"
          % (i_e - i_s, nnz_loc, bc, MF.col_panels, np.round(ms, 4).tolist(), 8 * (MF.handle.col_range()[1] - MF.handle.col_range()[0]) / 2 ** 20))
    MF.close()


if __name__ == "__main__":
    main()
