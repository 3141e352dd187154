#!/bin/bash
# Round 2, third single-GPU pass: suite (C power layer at world size 1, compressed panels, ...), the giant-row slowdown found by
# the config-3 sweep, second round of kernel experiments (128-thread CTAs, x gathers without L1 allocation), bench lines,
# fp32 after the lanes-per-row change, ncu (full capture of the CSR kernel at HEAD + launch list of the bench command),
# the rest of config 3 (resume), compute-sanitizer.
#   /usr/local/graft/bin/gpurun --timeout 3000 -- 'bash tools/gpu_round2d.sh'
O=gpurun_out/r02d
mkdir -p $O
SECONDS=0
timeout 900 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? t=${SECONDS}s" >> $O/pytest_gpu.log
timeout 300 python tools/repro_giant_rows.py --sell > $O/giant_rows.txt 2>&1; echo "giant rc=$? t=${SECONDS}s" >> $O/giant_rows.txt
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 150 --csv --log-file $O/launches_giant_rows.csv python tools/repro_giant_rows.py > $O/ncu_giant.log 2>&1; echo "ncu giant rc=$? t=${SECONDS}s" >> $O/ncu_giant.log
timeout 600 python tools/sweep_exp.py > $O/sweep_exp.jsonl 2> $O/sweep_exp.err; echo "sweep_exp rc=$? t=${SECONDS}s" >> $O/sweep_exp.err
timeout 600 python bench.py --steps 20 --warmup 5 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$? t=${SECONDS}s" >> $O/bench_n1.err
timeout 300 python bench.py --workload power --steps 20 --warmup 5 > $O/bench_power_n1.json 2> $O/bench_power_n1.err; echo "power rc=$? t=${SECONDS}s" >> $O/bench_power_n1.err
timeout 300 python tools/run_f32.py > $O/f32_vs_f64.jsonl 2> $O/f32_vs_f64.err; echo "f32 rc=$? t=${SECONDS}s" >> $O/f32_vs_f64.err
LINE="4329604 4329604 10 3.3333 normal random 0.6 100 0.95 0.5 14"
timeout 400 ncu --set full --clock-control none --import-source on -k regex:csr_tile_kernel -s 3 -c 1 -o $O/prof_csr python bench.py --steps 3 --warmup 3 --no-cpu-baseline --no-sell --format csr --line "$LINE" > $O/ncu_full_csr.log 2>&1; echo "ncu full rc=$? t=${SECONDS}s" >> $O/ncu_full_csr.log
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file $O/launches_bench.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-sell > $O/ncu_list.log 2>&1; echo "ncu list rc=$? t=${SECONDS}s" >> $O/ncu_list.log
timeout 800 python tools/run_medium_sweep.py --dataset twins_range --stride 8 --max-seconds 650 --resume profiles/r02b_config3_twins_range_part1.jsonl > $O/twins_range_stride8_part2.jsonl 2> $O/twins_range.err; echo "twins rc=$? t=${SECONDS}s" >> $O/twins_range.err
bash tools/gpu_sanitize.sh > $O/sanitize_stdout.log 2>&1
echo "done t=${SECONDS}s"
