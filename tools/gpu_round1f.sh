#!/bin/bash
# GPU pass r01f: AUTO format + SELL kernel shapes: parity tests, sweep, N=1 bench line
O=gpurun_out/r01f
mkdir -p $O
SECONDS=0
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? t=${SECONDS}s" >> $O/pytest_gpu.log
timeout 900 python tools/sweep.py full > $O/sweep.txt 2>&1; echo "sweep rc=$? t=${SECONDS}s" >> $O/sweep.txt
timeout 900 python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$? t=${SECONDS}s" >> $O/bench_n1.err
echo "done t=${SECONDS}s"
