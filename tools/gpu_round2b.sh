#!/bin/bash
# Round 2, second GPU pass (only after gpu_round2a.sh showed the device generator bit-identical on the B200): BASELINE config 2,
# the medium-dataset sweep generated on the GPU -- every 4th line of the 16 190 (systematic sample, all four footprint
# classes and the whole parameter grid), CSR vs SELL-C-sigma, cold device times, parity per matrix.  One GPU, ~30 min.
#   /usr/local/graft/bin/gpurun --timeout 2400 -- 'bash tools/gpu_round2b.sh'
O=gpurun_out/r02b
mkdir -p $O
SECONDS=0
timeout 2100 python tools/run_medium_sweep.py --stride 4 --max-seconds 1800 > $O/medium_sweep_stride4.jsonl 2> $O/medium_sweep.err
echo "sweep rc=$? t=${SECONDS}s" >> $O/medium_sweep.err
tail -n 1 $O/medium_sweep_stride4.jsonl > $O/medium_sweep_summary.json
echo "done t=${SECONDS}s"
