#!/bin/bash
# first GPU pass: parity tests, probe, bench, launch list, one full capture
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm,clocks.max.mem --format=csv > gpurun_out/box.txt 2>&1
lscpu | head -20 >> gpurun_out/box.txt; free -g >> gpurun_out/box.txt; nproc >> gpurun_out/box.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 ./tools/hbm_probe > gpurun_out/hbm_probe.txt 2>&1
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?" >> gpurun_out/bench_n1.err
LINE="4329604 4329604 10 3.3333 normal random 0.6 100 0.95 0.5 14"
CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-sell --line"
timeout 600 $CMD "$LINE" > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 60 --csv --log-file gpurun_out/launches.csv $CMD "$LINE" > gpurun_out/ncu_list.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:csr_tile_kernel -s 3 -c 1 -o gpurun_out/prof_csr $CMD "$LINE" > gpurun_out/ncu_full.log 2>&1
echo done
