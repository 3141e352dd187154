#!/bin/bash
O=gpurun_out/r01r
mkdir -p $O
SECONDS=0
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? t=${SECONDS}s" >> $O/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$? t=${SECONDS}s" >> $O/bench_n1.err
timeout 1500 python tools/run_configs.py classes > $O/configs.jsonl 2> $O/configs.txt; echo "configs rc=$? t=${SECONDS}s" >> $O/configs.txt
python - > $O/l2_probe.txt 2>&1 <<'PY'
import sys, numpy as np, torch
sys.path.insert(0, '.')
import spmv_research_b200 as S
from spmv_research_b200 import amg, harness as H
for ln in ("14713889 14713889 10 3.3333 normal random 0.6 0 1.9 0.5 14", "18448383 18448383 5 1.6667 normal random 0.6 10000 1.4 0.05 14"):
    p = amg.MatrixParams.parse(ln); rp, col, val = amg.generate(p); m, n, nnz = p.rows, p.cols, int(rp[-1])
    dx = torch.ones(n, dtype=torch.float64, device='cuda'); dy = torch.zeros(m, dtype=torch.float64, device='cuda')
    for fmt, fn in ((S.CSR_ADAPTIVE, 'csr'), (S.SELL_C_SIGMA, 'sell')):
        for mb, pan in ((4000, 1), (56, 1), (56, 2)):
            MF = S.csr_to_format(rp, col, val, m, n, nnz, format=fmt, col_panels=pan)
            MF.handle.time_device(dx, dy, 3, 256 << 20)
            ms = float(np.median(MF.handle.time_device(dx, dy, 9, 256 << 20)))
            print(ln[:50], fn, 'x_l2_mb', mb, 'panels', pan, '%.4f ms' % ms, flush=True)
            MF.close()
PY
echo "done t=${SECONDS}s"
