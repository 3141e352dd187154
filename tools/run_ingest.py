"""Input-side timings on the GPU box: generator pass 2 host vs device, COO -> CSR oracle (CPU) vs device,
each checked bit for bit before it is timed.  One JSON line per case.
  python tools/run_ingest.py [--small]"""
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    import torch
    from spmv_research_b200 import amg, capi
    from oracle import oracle as O          # checker only
    small = "--small" in sys.argv
    dev = "cuda:0"
    st = torch.cuda.current_stream().cuda_stream
    lines = ["200000 200000 10 5 normal random 0.3 0 0.5 0.5 14"] if small else [
        "1000000 1000000 10 5 normal random 0.3 0 0.5 0.5 14",                       # config 1
        "4194304 4194304 60 20 normal random 0.3 0 0.5 0.5 14",                      # one config-5 shard's worth
        "2000000 2000000 50 16.6667 gamma random 0.6 10000 0.05 0.05 14",            # giant first row (in-HBM sort)
        "2000000 2000000 50 150 gamma random 0.6 0 0.05 0.05 14",                    # heavy tail
    ]
    for line in lines:
        p = amg.MatrixParams.parse(line)
        t0 = time.time()
        rp, bw = amg.row_ptr(p)
        t_p1 = time.time() - t0
        t0 = time.time()
        col, val = amg.fill(p, rp, bw, 0, p.rows)
        t_host = time.time() - t0
        nnz = int(rp[-1])
        d_rp = torch.from_numpy(rp).to(dev)
        d_col = torch.empty(nnz, dtype=torch.int32, device=dev)
        d_val = torch.empty(nnz, dtype=torch.float64, device=dev)
        ts = []
        for _ in range(3):
            torch.cuda.synchronize()
            t0 = time.time()
            amg.fill_device(p, d_rp, bw, 0, p.rows, d_col, d_val, st)       # synchronises the stream itself
            ts.append(time.time() - t0)
        same = bool(np.array_equal(d_col.cpu().numpy(), col) and np.array_equal(d_val.cpu().numpy(), val))
        print(json.dumps({"what": "generator pass 2", "line": line, "nnz": nnz, "pass1_host_s": round(t_p1, 3), "host_s": round(t_host, 3),
                          "host_threads": os.cpu_count(), "device_s": [round(t, 4) for t in ts], "bit_identical": same}), flush=True)
        # COO -> CSR: the same matrix as shuffled COO
        if nnz <= 300_000_000:
            rng = np.random.default_rng(1)
            perm = rng.permutation(nnz)
            row = np.repeat(np.arange(p.rows, dtype=np.int32), np.diff(rp))[perm]
            c2, v2 = col[perm], val[perm]
            t0 = time.time()
            ref = O.coo_to_csr(row, c2, v2, p.rows, p.cols) if nnz <= 50_000_000 else None
            t_cpu = time.time() - t0
            d_r, d_c, d_v = (torch.from_numpy(a).to(dev) for a in (row, c2, v2))
            o_rp = torch.empty(p.rows + 1, dtype=torch.int32, device=dev)
            ts = []
            for _ in range(3):
                torch.cuda.synchronize()
                t0 = time.time()
                dups = capi.coo_to_csr_device(d_r, d_c, d_v, p.rows, p.cols, nnz, o_rp, d_col, d_val, st)
                ts.append(time.time() - t0)
            same = bool(np.array_equal(o_rp.cpu().numpy(), rp) and np.array_equal(d_col.cpu().numpy(), col) and np.array_equal(d_val.cpu().numpy(), val))
            print(json.dumps({"what": "coo_to_csr", "line": line, "nnz": nnz, "oracle_cpu_s": round(t_cpu, 3) if ref is not None else None,
                              "device_s": [round(t, 4) for t in ts], "duplicates": dups, "equals_generator_csr": same}), flush=True)
        del d_rp, d_col, d_val


if __name__ == "__main__":
    main()
