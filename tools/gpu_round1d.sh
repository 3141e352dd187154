#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
python - > gpurun_out/ref_harness.txt 2>&1 <<'PY'
import sys, json
sys.path.insert(0, '.')
from spmv_research_b200 import refharness as RH
line = "1000000 1000000 10 5 normal random 0.3 0 0.5 0.5 14".split()
for exe, gpu, env in (("spmv_csr_d.exe", False, None), ("spmv_b200_csr_d.exe", True, None), ("spmv_b200_sell_d.exe", True, None),
                      ("spmv_b200_csr_d.exe", True, {"B200SPMV_TRANSFER": "always"})):
    r = RH.run(exe, line, gpu_kernel=gpu, threads=16, extra_env=env)
    nnz = 10038177
    print(exe, env, "rc", r["returncode"], "loops", r.get("loops"), "median_s", r.get("time_median"), "GFLOPs(2nnz/t)", 2*nnz/r["time_median"]*1e-9 if r.get("time_median") else None)
    print("   csv:", ",".join(r.get("csv", [])))
PY
echo done
