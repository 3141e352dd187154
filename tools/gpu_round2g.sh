#!/bin/bash
# Round 2, sixth single-GPU pass: ncu --set full of the config-5 shard kernel without / with column panels (DRAM traffic vs
# compulsory bytes: the N = 8 limiter of round 1), of the SELL kernel at HEAD, then every 4th line of config 2 again at HEAD
# (single-launch products, conditional split-row fold).
#   /usr/local/graft/bin/gpurun --timeout 1800 -- 'bash tools/gpu_round2g.sh'
O=gpurun_out/r02g
mkdir -p $O
SECONDS=0
timeout 300 python tools/profile_c5_shard.py --panels 1 > $O/c5shard_p1.txt 2>&1
timeout 300 python tools/profile_c5_shard.py --panels 2 > $O/c5shard_p2.txt 2>&1; echo "shard rc=$? t=${SECONDS}s" >> $O/c5shard_p2.txt
timeout 400 ncu --set full --clock-control none --import-source on -k regex:csr_tile_kernel -s 2 -c 1 -o $O/prof_c5shard_p1 python tools/profile_c5_shard.py --panels 1 --iters 3 > $O/ncu_c5_p1.log 2>&1; echo "ncu p1 rc=$? t=${SECONDS}s" >> $O/ncu_c5_p1.log
timeout 400 ncu --set full --clock-control none --import-source on -k regex:csr_tile_kernel -s 4 -c 2 -o $O/prof_c5shard_p2 python tools/profile_c5_shard.py --panels 2 --iters 3 > $O/ncu_c5_p2.log 2>&1; echo "ncu p2 rc=$? t=${SECONDS}s" >> $O/ncu_c5_p2.log
LINE="2200290 2200290 20 6.6667 normal random 0.3 1000 1.4 0.95 14"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:sell_kernel -s 3 -c 1 -o $O/prof_sell python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-sell --format sell --line "$LINE" > $O/ncu_full_sell.log 2>&1; echo "ncu sell rc=$? t=${SECONDS}s" >> $O/ncu_full_sell.log
timeout 1000 python tools/run_medium_sweep.py --stride 4 --max-seconds 900 > $O/medium_sweep_stride4_head.jsonl 2> $O/medium_sweep.err; echo "sweep rc=$? t=${SECONDS}s" >> $O/medium_sweep.err
echo "done t=${SECONDS}s"
