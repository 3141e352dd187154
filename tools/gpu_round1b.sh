#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 3 > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench rc=$?" >> gpurun_out/bench_n1.err
KEXP_VARIANTS="lib csr auto,lib sell C32" timeout 600 ./tools/kexp "1000000 1000000 10 5 normal random 0.3 0 0.5 0.5 14" "8388607 8388607 50 150 gamma random 0.6 0 0.05 0.05 14" "8388607 8388607 50 16.6667 gamma random 0.6 10000 0.05 0.05 14" "89418 89418 500 166.667 normal random 0.3 100 0.5 0.5 14" "4194304 4194304 60 20 normal random 0.3 0 0.5 0.5 14" > gpurun_out/kexp2.txt 2>&1
echo done
