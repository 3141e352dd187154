#!/bin/bash
mkdir -p gpurun_out
timeout 600 ./tools/kexp "4329604 4329604 10 3.3333 normal random 0.6 100 0.95 0.5 14" "4329604 4329604 10 3.3333 normal random 0.05 0 0.05 0.05 14" "888859 888859 50 16.6667 normal random 0.6 0 0.5 0.5 14" "5242879 5242879 5 1.6667 normal random 0.3 0 0.5 0.5 14" "445906 445906 100 33.3333 normal random 0.05 0 0.95 0.95 14" > gpurun_out/kexp.txt 2>&1
