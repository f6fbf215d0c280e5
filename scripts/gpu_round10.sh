#!/bin/bash
set -u
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_golden.py tests/test_gpu_parity.py tests/test_gpu_operators.py -m gpu -q -x > gpurun_out/pytest_load.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/pytest_load.log
timeout 300 python bench.py --steps 5 --warmup 3 --krylov-iters 20 --no-mg 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('step ms', d['ms_per_step'], 'e2e', d['e2e'])"
