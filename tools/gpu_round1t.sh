#!/bin/bash
# final pass of round 1: parity tests, N=1 bench line, reference arm, launch list of the bench command, full captures of the
# two kernels at HEAD, the reference's unmodified harness driving the plugin (csr / sell / auto, transfer first / always)
O=gpurun_out/r01t
mkdir -p $O
SECONDS=0
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? t=${SECONDS}s" >> $O/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$? t=${SECONDS}s" >> $O/bench_n1.err
timeout 600 python bench.py --workload power --steps 10 --warmup 3 --no-graph > $O/bench_power_n1.json 2> $O/bench_power_n1.err; echo "power rc=$? t=${SECONDS}s" >> $O/bench_power_n1.err
timeout 900 python bench.py --impl reference --steps 5 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$? t=${SECONDS}s" >> $O/bench_ref.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sell > $O/ncu_list.log 2>&1
LINE="4329604 4329604 10 3.3333 normal random 0.6 100 0.95 0.5 14"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:csr_tile_kernel -s 3 -c 1 -o $O/prof_csr python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-sell --format csr --line "$LINE" > $O/ncu_full_csr.log 2>&1
python - > $O/ref_harness.txt 2>&1 <<'PY'
import sys
sys.path.insert(0, '.')
from spmv_research_b200 import refharness as RH
line = "1000000 1000000 10 5 normal random 0.3 0 0.5 0.5 14".split()
nnz = 10038177
for exe, gpu, env in (("spmv_csr_d.exe", False, None), ("spmv_b200_csr_d.exe", True, None), ("spmv_b200_sell_d.exe", True, None), ("spmv_b200_auto_d.exe", True, None),
                      ("spmv_b200_auto_d.exe", True, {"B200SPMV_TRANSFER": "always"}), ("spmv_b200_auto_d.exe", True, {"B200SPMV_TRANSFER": "always", "B200SPMV_PIN": "0"})):
    r = RH.run(exe, line, gpu_kernel=gpu, threads=16, extra_env=env)
    print(exe, env, "rc", r["returncode"], "loops", r.get("loops"), "median_s", r.get("time_median"), "GFLOPs(2nnz/t)", 2 * nnz / r["time_median"] * 1e-9 if r.get("time_median") else None)
    print("   csv:", ",".join(r.get("csv", [])))
PY
echo "done t=${SECONDS}s"
