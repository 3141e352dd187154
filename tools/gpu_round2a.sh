#!/bin/bash
# Round 2, first GPU pass: (1) the suite, (2) the tests written after round 1's GPU budget ran out (host_pipeline, device
# generator, COO -> CSR, fp32 product + _f harness), each file on its own so that one failure does not hide the others, (3) bench default vs
# --host-pipeline (e2e), (4) the N=1 power member with the host and the device generator (generate_s), (5) ingest timings,
# (6) launch list of the device generator + COO conversion.  One GPU.
#   /usr/local/graft/bin/gpurun --timeout 2400 -- 'bash tools/gpu_round2a.sh'
O=gpurun_out/r02a
mkdir -p $O
SECONDS=0
timeout 900 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? t=${SECONDS}s" >> $O/pytest_gpu.log
for t in host_pipeline amg_device coo_device f32_device plain_sell; do
	B200SPMV_TEST_NEW=1 timeout 600 python -m pytest tests/test_zz_$t.py -m gpu -q > $O/pytest_new_$t.log 2>&1; echo "pytest $t rc=$? t=${SECONDS}s" >> $O/pytest_new_$t.log
done
timeout 600 python tools/run_f32.py > $O/f32_vs_f64.jsonl 2> $O/f32_vs_f64.err; echo "f32 rc=$? t=${SECONDS}s" >> $O/f32_vs_f64.err
timeout 600 python tools/run_ingest.py > $O/ingest.jsonl 2> $O/ingest.err; echo "ingest rc=$? t=${SECONDS}s" >> $O/ingest.err
timeout 900 python bench.py --steps 20 --warmup 3 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$? t=${SECONDS}s" >> $O/bench_n1.err
timeout 900 python bench.py --steps 20 --warmup 3 --host-pipeline --no-cpu-baseline > $O/bench_n1_pipeline.json 2> $O/bench_n1_pipeline.err; echo "bench pipeline rc=$? t=${SECONDS}s" >> $O/bench_n1_pipeline.err
timeout 600 python bench.py --workload power --steps 10 --warmup 3 --no-graph > $O/bench_power_n1.json 2> $O/bench_power_n1.err; echo "power rc=$? t=${SECONDS}s" >> $O/bench_power_n1.err
timeout 600 python bench.py --workload power --steps 10 --warmup 3 --no-graph --device-generator > $O/bench_power_n1_devgen.json 2> $O/bench_power_n1_devgen.err; echo "power devgen rc=$? t=${SECONDS}s" >> $O/bench_power_n1_devgen.err
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 200 --csv --log-file $O/launches_ingest.csv python tools/run_ingest.py --small > $O/ncu_ingest.log 2>&1; echo "ncu rc=$? t=${SECONDS}s" >> $O/ncu_ingest.log
echo "done t=${SECONDS}s"
