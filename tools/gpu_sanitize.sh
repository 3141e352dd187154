#!/bin/bash
# compute-sanitizer over the edge-case set: memcheck (out-of-bounds / misaligned accesses, leaks of device
# memory are not tracked) and racecheck (shared-memory hazards: the tile kernel's product round trip, the bitonic sort
# phases of the conversion, the reductions).  One GPU, ~10 min.  Logs: gpurun_out/sanitize/.
#   /usr/local/graft/bin/gpurun --timeout 1500 -- 'bash tools/gpu_sanitize.sh'
O=gpurun_out/sanitize
mkdir -p $O
SECONDS=0
CS=/usr/local/cuda/bin/compute-sanitizer
timeout 900 $CS --tool memcheck --error-exitcode 9 --print-limit 20 python tools/run_sanitize_cases.py > $O/memcheck.log 2>&1; echo "memcheck rc=$? t=${SECONDS}s" >> $O/memcheck.log
timeout 900 $CS --tool racecheck --racecheck-report analysis --error-exitcode 9 --print-limit 20 python tools/run_sanitize_cases.py > $O/racecheck.log 2>&1; echo "racecheck rc=$? t=${SECONDS}s" >> $O/racecheck.log
grep -E "ERROR SUMMARY|RACECHECK SUMMARY|rc=|wrong result" $O/memcheck.log $O/racecheck.log > $O/summary.txt
echo "done t=${SECONDS}s"
