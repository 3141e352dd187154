"""csr_to_format cost per format on the bench sample (+ config 1, G2): device time of the format construction (b200spmv_stats
convert_ms: upload of the CSR arrays from pinned-or-not host memory included, so the host-array create is also run from DEVICE
arrays, where it is the on-GPU conversion alone).  One JSON line per matrix.   python tools/run_convert_times.py > out.jsonl"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch  # noqa: E402
import spmv_research_b200 as S  # noqa: E402
from spmv_research_b200 import amg, formats, harness as H  # noqa: E402


def main():
    dev = torch.device("cuda:0")
    plist = [H.CONFIG1] + H.medium_sample() + ([H.CONFIG4_G2] if "--g2" in sys.argv else [])
    for p in plist:
        rp, d_rp, d_col, d_val = amg.generate_device(p)
        m, n, nnz = p.rows, p.cols, int(rp[-1])
        rec = {"matrix": p.line(), "nnz": nnz}
        for name, kw in (("csr", dict(format=S.CSR_ADAPTIVE)), ("sell_16384", dict(format=S.SELL_C_SIGMA)),
                         ("sell_256", dict(format=S.SELL_C_SIGMA, sell_sigma=256)), ("auto", dict(format=S.AUTO))):
            ts = []
            for _ in range(3):
                MF = formats.csr_to_format_device(d_rp, d_col, d_val, m, n, nnz, **kw)
                ts.append(MF.convert_ms)
                MF.close()
            rec[name + "_convert_ms"] = round(float(np.median(ts)), 3)
        print(json.dumps(rec), flush=True)
        del d_rp, d_col, d_val
        torch.cuda.empty_cache()


if __name__ == "__main__":
    main()
