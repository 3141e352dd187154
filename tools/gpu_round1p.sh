#!/bin/bash
O=gpurun_out/r01p
mkdir -p $O
SECONDS=0
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? t=${SECONDS}s" >> $O/pytest_gpu.log
SWEEP_PANELS=1 timeout 1500 python tools/sweep_l2.py > $O/sweep_panels.txt 2>&1; echo "sweep rc=$? t=${SECONDS}s" >> $O/sweep_panels.txt
echo "done t=${SECONDS}s"
