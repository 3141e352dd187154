#!/bin/bash
# What the driver runs at round end, on one GPU: the GPU suite (-x), smoke(), the default bench line, the reference arm.
O=gpurun_out/final
mkdir -p $O
SECONDS=0
timeout 600 python -m pytest tests -m gpu -x -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? t=${SECONDS}s" >> $O/pytest_gpu.log
timeout 200 python -c "import __graft_entry__ as g; g.smoke()" > $O/smoke.log 2>&1; echo "smoke rc=$? t=${SECONDS}s" >> $O/smoke.log
timeout 400 python bench.py --gpus 1 --steps 20 --warmup 5 > $O/bench_n1.json 2> $O/bench_n1.err; echo "bench rc=$? t=${SECONDS}s" >> $O/bench_n1.err
timeout 400 python bench.py --impl reference --gpus 1 --steps 20 --warmup 5 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$? t=${SECONDS}s" >> $O/bench_ref.err
echo "done t=${SECONDS}s"
