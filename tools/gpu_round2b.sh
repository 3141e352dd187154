#!/bin/bash
# Round 2, second GPU pass: the whole suite un-gated, then BASELINE config 2 -- the medium-dataset sweep generated on the GPU,
# every 4th line of the 16 190 (systematic sample: all four footprint classes, the whole parameter grid), CSR vs
# SELL-C-sigma, cold device times, parity per matrix -- then config 3's full parameter file (the +-30 % expansion of the 52
# twins), every 8th line.  One GPU.  An interrupted sweep continues with --resume <earlier jsonl>.
#   /usr/local/graft/bin/gpurun --timeout 3300 -- 'bash tools/gpu_round2b.sh'
O=gpurun_out/r02b
mkdir -p $O
SECONDS=0
timeout 900 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? t=${SECONDS}s" >> $O/pytest_gpu.log
timeout 2000 python tools/run_medium_sweep.py --stride 4 --max-seconds 1700 > $O/medium_sweep_stride4.jsonl 2> $O/medium_sweep.err
echo "sweep rc=$? t=${SECONDS}s" >> $O/medium_sweep.err
tail -n 1 $O/medium_sweep_stride4.jsonl > $O/medium_sweep_summary.json
timeout 900 python tools/run_medium_sweep.py --dataset twins_range --stride 8 --max-seconds 700 > $O/twins_range_stride8.jsonl 2> $O/twins_range.err
echo "twins_range rc=$? t=${SECONDS}s" >> $O/twins_range.err
tail -n 1 $O/twins_range_stride8.jsonl > $O/twins_range_summary.json
echo "done t=${SECONDS}s"
