#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
KEXP_VARIANTS="lib csr auto,lib sell C32" timeout 600 ./tools/kexp "1000000 1000000 10 5 normal random 0.3 0 0.5 0.5 14" "8388607 8388607 50 150 gamma random 0.6 0 0.05 0.05 14" "8388607 8388607 50 16.6667 gamma random 0.6 10000 0.05 0.05 14" "89418 89418 500 166.667 normal random 0.3 100 0.5 0.5 14" "4194304 4194304 60 20 normal random 0.3 0 0.5 0.5 14" "4329604 4329604 10 3.3333 normal random 0.6 100 0.95 0.5 14" "888859 888859 50 16.6667 normal random 0.3 10000 1.9 0.05 14" "2200290 2200290 20 6.6667 normal random 0.3 1000 1.4 0.95 14" > gpurun_out/kexp2.txt 2>&1
echo done
