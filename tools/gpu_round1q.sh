#!/bin/bash
O=gpurun_out/r01q
mkdir -p $O
SECONDS=0
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? t=${SECONDS}s" >> $O/pytest_gpu.log
timeout 1500 python tools/run_configs.py config5 classes > $O/configs.jsonl 2> $O/configs.txt; echo "configs rc=$? t=${SECONDS}s" >> $O/configs.txt
timeout 900 python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$? t=${SECONDS}s" >> $O/bench_n1.err
echo "done t=${SECONDS}s"
