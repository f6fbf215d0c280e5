#!/bin/bash
set -u
mkdir -p gpurun_out
for v in 8 88 4 48; do
FEMO_SPMV=$v timeout 300 python bench.py --steps 3 --warmup 3 --krylov-iters 100 --no-mg 2>/dev/null | python -c "
import json,sys
d=json.loads(sys.stdin.read().strip().splitlines()[-1]); print('variant $v s/iter', d['solve']['s_per_krylov_iter'], 'e2e ms', d['e2e']['ms_per_step'])"
done
FEMO_SPMV=48 timeout 300 python -m pytest tests/test_gpu_golden.py tests/test_gpu_multigrid.py -m gpu -q -x 2>&1 | tail -n 1
