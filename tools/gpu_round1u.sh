#!/bin/bash
O=gpurun_out/r01u
mkdir -p $O
SECONDS=0
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? t=${SECONDS}s" >> $O/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$? t=${SECONDS}s" >> $O/bench_n1.err
timeout 600 python bench.py --workload power --steps 10 --warmup 3 --no-graph > $O/bench_power_n1.json 2> $O/bench_power_n1.err; echo "power rc=$? t=${SECONDS}s" >> $O/bench_power_n1.err
echo "done t=${SECONDS}s"
