#!/bin/bash
# multi-GPU bench line under torchrun, tight timeouts (a hung rank keeps N GPUs leased)
mkdir -p gpurun_out
N=${1:-2}
nvidia-smi topo -m > gpurun_out/topo_n$N.txt 2>&1
SECONDS=0
timeout ${TMO:-300} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "rc=$? wall=${SECONDS}s" >> gpurun_out/bench_n$N.err
echo done
