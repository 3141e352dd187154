#!/bin/bash
# Round 2, kernel experiments on the CSR row reduction (the stage that costs ~30 % over the stream + gather floor):
# the shipped library against builds with LDS.128 pair reads, >= 2 / 4 / 8 lanes per row, the conflict-free row -> thread table
# (CSR_EXP_PERM) and the per-row start rotation (CSR_EXP_ROT) -- csr_kernel.cu switches, csr_perm.h.
# Build the experimental libraries HERE first (they travel with the snapshot, git-ignored):
#   python -c "from spmv_research_b200 import build as B; B.build_experiments()"
#   /usr/local/graft/bin/gpurun --timeout 1500 -- 'bash tools/gpu_round2c.sh'
O=gpurun_out/r02c
mkdir -p $O
SECONDS=0
timeout 1400 python tools/sweep_exp.py > $O/sweep_exp.jsonl 2> $O/sweep_exp.err; echo "sweep_exp rc=$? t=${SECONDS}s" >> $O/sweep_exp.err
grep -E "summary|drift" $O/sweep_exp.jsonl > $O/sweep_exp_summary.jsonl
echo "done t=${SECONDS}s"
