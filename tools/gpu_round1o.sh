#!/bin/bash
# ncu --set full captures: kernel family x config (SURVEY 8d evidence): config 1 (CSR), G1 (SELL), G2 (CSR), config-5 shard (CSR, SELL)
O=gpurun_out/r01o
mkdir -p $O
SECONDS=0
cap() { # name format kernel-regex line
  CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-sell --format $2 --line"
  timeout 300 $CMD "$4" > $O/plain_$1.log 2>&1 &&
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:$3 -s 3 -c 1 -o $O/prof_$1 $CMD "$4" > $O/ncu_$1.log 2>&1
  echo "$1 rc=$? t=${SECONDS}s" >> $O/log.txt
}
cap config1_csr csr csr_tile_kernel "1000000 1000000 10 5 normal random 0.3 0 0.5 0.5 14"
cap g1_sell sell sell_kernel "8388607 8388607 50 150 gamma random 0.6 0 0.05 0.05 14"
cap g2_csr csr csr_tile_kernel "8388607 8388607 50 16.6667 gamma random 0.6 10000 0.05 0.05 14"
cap c5shard_csr csr csr_tile_kernel "4194304 33554432 60 20 normal random 0.3 0 0.5 0.5 14"
cap c5shard_sell sell sell_kernel "4194304 33554432 60 20 normal random 0.3 0 0.5 0.5 14"
echo "done t=${SECONDS}s"
