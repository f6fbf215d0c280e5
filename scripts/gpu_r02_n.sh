#!/bin/bash
# 1-GPU session: fp32-smoother comparison of the multigrid solve, multigrid parity tests, ncu --set full of the assembly kernel
set -u
mkdir -p gpurun_out
T=${TAG:-r02n}
for f in 1 0; do
  FEMO_MG_FP32=$f FEMO_VERBOSE=1 timeout 600 python scripts/mg_bench.py 51 5 3 1e-10 > gpurun_out/${T}_mg_fp32_$f.log 2>&1
  echo "FEMO_MG_FP32=$f rc=$?"; grep -v "tile plan\|mg level\|direct solver" gpurun_out/${T}_mg_fp32_$f.log | tail -n 3 | cut -c1-700
done
timeout 900 python -m pytest tests/test_gpu_multigrid.py tests/test_gpu_direct.py tests/test_gpu_dynamic.py -m gpu -q --tb=short -s > gpurun_out/${T}_pytest_mg.log 2>&1; echo "pytest mg rc=$?"
grep -E "passed|failed|V-cycle|wing box x" gpurun_out/${T}_pytest_mg.log | cut -c1-250 | tail -n 12
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_asm_tiles_ws -s 3 -c 1 -f -o gpurun_out/${T}_asm_tiles python scripts/asm_experiments.py --masks 0 > gpurun_out/${T}_ncu_asm.log 2>&1; echo "ncu asm rc=$?"
ls -la gpurun_out/${T}_asm_tiles.ncu-rep
