#!/bin/bash
O=gpurun_out/r01h
mkdir -p $O
SECONDS=0
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? t=${SECONDS}s" >> $O/pytest_gpu.log
timeout 1500 python tools/run_configs.py config1 twins g1 g2 > $O/configs.jsonl 2> $O/configs.txt; echo "configs rc=$? t=${SECONDS}s" >> $O/configs.txt
echo "done t=${SECONDS}s"
