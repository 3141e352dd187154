#!/bin/bash
# Round 2, fifth single-GPU pass: the suite with the folded split-row / cut-slice finish and the range launches of the host path,
# bench lines, every 16th line of config 2 again (effect of the single-launch products per footprint class), launch list.
#   /usr/local/graft/bin/gpurun --timeout 1500 -- 'bash tools/gpu_round2f.sh'
O=gpurun_out/r02f
mkdir -p $O
SECONDS=0
timeout 900 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? t=${SECONDS}s" >> $O/pytest_gpu.log
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$? t=${SECONDS}s" >> $O/bench_n1.err
timeout 300 python bench.py --workload power --steps 20 --warmup 5 > $O/bench_power_n1.json 2> $O/bench_power_n1.err; echo "power rc=$? t=${SECONDS}s" >> $O/bench_power_n1.err
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$? t=${SECONDS}s" >> $O/smoke.log
timeout 600 python tools/run_medium_sweep.py --stride 16 --offset 3 --max-seconds 400 --check-every 25 > $O/medium_sweep_stride16.jsonl 2> $O/medium_sweep.err; echo "sweep rc=$? t=${SECONDS}s" >> $O/medium_sweep.err
timeout 300 python tools/run_configs.py config1 g1 g2 > $O/configs_1_4.jsonl 2> $O/configs_1_4.txt; echo "configs rc=$? t=${SECONDS}s" >> $O/configs_1_4.txt
timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file $O/launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sell > $O/ncu_list.log 2>&1; echo "ncu list rc=$? t=${SECONDS}s" >> $O/ncu_list.log
echo "done t=${SECONDS}s"
