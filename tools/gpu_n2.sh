#!/bin/bash
# N GPUs of one box (default 2): the multi-GPU layer under the C-ABI -- pytest (world-2 cases of tests/test_zz_power_c.py: per-step
# oracle parity on every rank, received band rows bit-equal, C host program with forked ranks), then the bench line under torchrun.
# Tight timeouts: a hung rank keeps N GPUs leased.
#   /usr/local/graft/bin/gpurun --gpus 2 --timeout 900 -- 'bash tools/gpu_n2.sh 2'
N=${1:-2}
O=gpurun_out/r02n$N
mkdir -p $O
nvidia-smi topo -m > $O/topo.txt 2>&1
SECONDS=0
if [ "$N" = "2" ]; then
	timeout 900 python -m pytest tests -m gpu -q > $O/pytest_gpu.log 2>&1; echo "pytest rc=$? t=${SECONDS}s" >> $O/pytest_gpu.log
	timeout 200 tests/c/power_test 2 400000 > $O/power_test_c.log 2>&1; echo "power_test rc=$? t=${SECONDS}s" >> $O/power_test_c.log
fi
timeout ${TMO:-420} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 20 --warmup 3 > $O/bench.json 2> $O/bench.err; echo "bench rc=$? t=${SECONDS}s" >> $O/bench.err
if [ "$N" = "2" ]; then
	# the reference arm keeps N GPUs leased while the host cores work: only at N = 2 here, with a short budget (the driver runs it at every N)
	B200_REF_BUDGET_S=60 timeout ${TMO:-420} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus $N --steps 3 --warmup 1 > $O/bench_ref.json 2> $O/bench_ref.err; echo "ref rc=$? t=${SECONDS}s" >> $O/bench_ref.err
fi
echo "done t=${SECONDS}s"
