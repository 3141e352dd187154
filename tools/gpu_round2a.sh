#!/bin/bash
# Round 2, first GPU pass: (1) the suite, (2) the tests written after round 1's GPU budget ran out (host_pipeline),
# (3) bench default vs --host-pipeline (e2e), (4) the N=1 power member.  One GPU.
#   /usr/local/graft/bin/gpurun --timeout 1500 -- 'bash tools/gpu_round2a.sh'
O=gpurun_out/r02a
mkdir -p $O
SECONDS=0
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? t=${SECONDS}s" >> $O/pytest_gpu.log
B200SPMV_TEST_NEW=1 timeout 600 python -m pytest tests/test_zz_host_pipeline.py -m gpu -q > $O/pytest_new.log 2>&1; echo "pytest new rc=$? t=${SECONDS}s" >> $O/pytest_new.log
timeout 900 python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$? t=${SECONDS}s" >> $O/bench_n1.err
timeout 900 python bench.py --steps 20 --warmup 3 --host-pipeline --no-cpu-baseline > $O/bench_n1_pipeline.json 2> $O/bench_n1_pipeline.err; echo "bench pipeline rc=$? t=${SECONDS}s" >> $O/bench_n1_pipeline.err
timeout 600 python bench.py --workload power --steps 10 --warmup 3 --no-graph > $O/bench_power_n1.json 2> $O/bench_power_n1.err; echo "power rc=$? t=${SECONDS}s" >> $O/bench_power_n1.err
echo "done t=${SECONDS}s"
