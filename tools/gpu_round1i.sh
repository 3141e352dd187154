#!/bin/bash
# ncu on the two slow twins (mawi: one 16M-nnz row + 18M rows of ~1 nnz; kmer: 55M rows x 2, x band > L2)
O=gpurun_out/r01i
mkdir -p $O
M1="18571154 18571154 2.0483552072 3805.8118835271 normal random 0.0090474297 8006372.0851891888 0.8839521855 0.8874492450 14"
M2="55042369 55042369 2.1295885720 0.6731069663 normal random 0.3370633097 17.3133965463 0.0573248727 0.1163647359 14"
i=0
for L in "$M1" "$M2"; do
  i=$((i+1))
  CMD="python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-sell --format csr --line"
  timeout 300 $CMD "$L" > $O/plain_$i.log 2>&1 &&
  timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 40 --csv --log-file $O/launches_$i.csv $CMD "$L" > $O/ncu_list_$i.log 2>&1
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:csr_tile_kernel -s 3 -c 1 -o $O/prof_csr_$i $CMD "$L" > $O/ncu_full_$i.log 2>&1
done
echo done
