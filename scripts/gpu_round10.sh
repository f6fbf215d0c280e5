#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 120 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "tiled_assembly_many" > gpurun_out/pytest_bal.log 2>&1; rc=$?; echo "pytest tiled rc=$rc"; tail -n 2 gpurun_out/pytest_bal.log
if [ $rc -ne 0 ]; then exit 1; fi
timeout 300 python bench.py --steps 10 --warmup 3 --krylov-iters 5 --no-mg 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('step ms', d['ms_per_step'], 'asm ms', d['roofline']['avg_launch_ms'], 'frac', d['roofline']['frac'], 'e2e', d['e2e']['ms_per_step'])"
