#!/bin/bash
# GPU pass r01e (HEAD): parity tests, N=1 bench line, variant sweep, launch list + full captures (CSR and SELL)
O=gpurun_out/r01e
mkdir -p $O
SECONDS=0
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? t=${SECONDS}s" >> $O/pytest_gpu.log
timeout 900 python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$? t=${SECONDS}s" >> $O/bench_n1.err
timeout 900 python tools/sweep.py full > $O/sweep.txt 2>&1; echo "sweep rc=$? t=${SECONDS}s" >> $O/sweep.txt
LINE="4329604 4329604 10 3.3333 normal random 0.6 100 0.95 0.5 14"
for F in csr sell; do
  CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-sell --format $F --line"
  timeout 300 $CMD "$LINE" > $O/plain_$F.log 2>&1 &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 80 --csv --log-file $O/launches_$F.csv $CMD "$LINE" > $O/ncu_list_$F.log 2>&1
done
timeout 600 ncu --set full --clock-control none --import-source on -k regex:csr_tile_kernel -s 3 -c 1 -o $O/prof_csr python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-sell --format csr --line "$LINE" > $O/ncu_full_csr.log 2>&1
timeout 600 ncu --set full --clock-control none --import-source on -k regex:sell_kernel -s 3 -c 1 -o $O/prof_sell python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-sell --format sell --line "$LINE" > $O/ncu_full_sell.log 2>&1
echo "done t=${SECONDS}s"
