#!/bin/bash
# 1-GPU session: tiled-kernel writer variants -- parity of the default build, then kernel time of each build at config 3
set -u
mkdir -p gpurun_out
T=${TAG:-r02k2}
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q --tb=short -k "tiled or halo_records or benchmark_size or jacobian_matrix or deterministic" > gpurun_out/${T}_pytest_tiles.log 2>&1; echo "pytest tiles rc=$?"
grep -E "passed|failed" gpurun_out/${T}_pytest_tiles.log | tail -n 3
for v in old u1 u1s; do
  FEMO_LIB_PATH=$PWD/femo_alpha_b200/libfemo_b200_$v.so timeout 300 python scripts/asm_experiments.py --masks 0,4 > gpurun_out/${T}_asm_$v.log 2>&1; echo "variant $v rc=$?"; grep "mask" gpurun_out/${T}_asm_$v.log
done
