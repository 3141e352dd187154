#!/bin/bash
O=gpurun_out/r01k
mkdir -p $O
SECONDS=0
timeout 1500 python tools/run_configs.py classes > $O/classes.jsonl 2> $O/classes.txt; echo "classes rc=$? t=${SECONDS}s" >> $O/classes.txt
timeout 1500 python tools/run_configs.py config5 > $O/config5.jsonl 2> $O/config5.txt; echo "config5 rc=$? t=${SECONDS}s" >> $O/config5.txt
echo "done t=${SECONDS}s"
