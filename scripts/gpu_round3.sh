#!/bin/bash
set -u
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi_L.txt
python -m pytest tests/test_gpu_dist.py tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_gpu2.log 2>&1; echo "pytest rc=$?"; tail -n 25 gpurun_out/pytest_gpu2.log
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 5 --warmup 3 --krylov-iters 50 > gpurun_out/bench_2gpu.json 2> gpurun_out/bench_2gpu.err; echo "bench2 rc=$?"; tail -n 2 gpurun_out/bench_2gpu.json; tail -n 5 gpurun_out/bench_2gpu.err
