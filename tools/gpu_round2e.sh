#!/bin/bash
# Round 2, fourth single-GPU pass: suite at HEAD (edge-case set through every path, long-row panel split, compressed panels),
# kmer twin panel sweep, configs 1 / 4 at HEAD, bench + its launch list (enough launches to reach the timed steps), then the
# REST of config 2: every line of the medium dataset not yet measured (resume from the stride-4 run).
#   /usr/local/graft/bin/gpurun --timeout 3400 -- 'bash tools/gpu_round2e.sh'
O=gpurun_out/r02e
mkdir -p $O
SECONDS=0
timeout 900 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? t=${SECONDS}s" >> $O/pytest_gpu.log
timeout 400 python tools/run_kmer_panels.py > $O/kmer_panels.jsonl 2> $O/kmer_panels.err; echo "kmer rc=$? t=${SECONDS}s" >> $O/kmer_panels.err
timeout 600 python tools/run_configs.py config1 g1 g2 > $O/configs_1_4.jsonl 2> $O/configs_1_4.txt; echo "configs rc=$? t=${SECONDS}s" >> $O/configs_1_4.txt
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$? t=${SECONDS}s" >> $O/bench_n1.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 2500 --csv --log-file $O/launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sell > $O/ncu_list.log 2>&1; echo "ncu list rc=$? t=${SECONDS}s" >> $O/ncu_list.log
gunzip -c profiles/r02b_config2_sweep_stride4.jsonl.gz > /tmp/stride4.jsonl
timeout 2400 python tools/run_medium_sweep.py --stride 1 --max-seconds 2250 --check-every 100 --resume /tmp/stride4.jsonl > $O/medium_sweep_rest.jsonl 2> $O/medium_sweep_rest.err; echo "sweep rc=$? t=${SECONDS}s" >> $O/medium_sweep_rest.err
echo "done t=${SECONDS}s"
