#!/bin/bash
# N-GPU bench line under torchrun (default exchange = NVSwitch multicast, band-limited), tight timeouts
mkdir -p gpurun_out/r01g
N=${1:-2}
O=gpurun_out/r01g
nvidia-smi topo -m > $O/topo_n$N.txt 2>&1
SECONDS=0
NCCL_DEBUG=WARN timeout ${TMO:-300} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 10 --warmup 3 > $O/bench_n$N.json 2> $O/bench_n$N.err; echo "rc=$? wall=${SECONDS}s" >> $O/bench_n$N.err
if [ "$2" = "p2p" ]; then
timeout ${TMO:-300} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus $N --steps 10 --warmup 3 --exchange p2p --no-nccl-leg > $O/bench_n${N}_p2p.json 2> $O/bench_n${N}_p2p.err; echo "rc=$? wall=${SECONDS}s" >> $O/bench_n${N}_p2p.err
timeout ${TMO:-300} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --gpus $N --steps 10 --warmup 3 --full-exchange --no-nccl-leg > $O/bench_n${N}_mcfull.json 2> $O/bench_n${N}_mcfull.err; echo "rc=$? wall=${SECONDS}s" >> $O/bench_n${N}_mcfull.err
fi
echo done
