#!/bin/bash
# Round 2: BASELINE config 3 widened to every 2nd line of the 3 637-line range file (resume from the every-8th runs), the mawi
# twin's giant-skew variants skipped (--max-skew, counted in the summary).  One GPU, <= 9 min.
O=gpurun_out/r02j
mkdir -p $O
SECONDS=0
timeout 560 python tools/run_medium_sweep.py --dataset twins_range --stride 2 --max-skew 1e6 --max-seconds 500 --check-every 40 \
	--resume profiles/r02b_config3_twins_range_part1.jsonl --resume profiles/r02d_config3_twins_range_part2.jsonl \
	> $O/twins_range_stride2.jsonl 2> $O/twins_range.err; echo "twins rc=$? t=${SECONDS}s" >> $O/twins_range.err
echo "done t=${SECONDS}s"
