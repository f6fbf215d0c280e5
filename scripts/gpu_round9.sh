#!/bin/bash
# multi-GPU: dist parity test + bench at N = $1
set -u
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L | head -8
if [ "$N" = "2" ]; then timeout 600 python -m pytest tests/test_gpu_dist.py -m gpu -x -q > gpurun_out/pytest_dist_n$N.log 2>&1; echo "pytest dist rc=$?"; tail -n 5 gpurun_out/pytest_dist_n$N.log; fi
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --steps 5 --warmup 3 --krylov-iters 50 > gpurun_out/bench_n$N.json 2> gpurun_out/bench_n$N.err; echo "bench N=$N rc=$?"; tail -n 1 gpurun_out/bench_n$N.json; grep -v "OMP_NUM\|^\*\*\*\|^$" gpurun_out/bench_n$N.err | tail -n 5
