#!/bin/bash
# 1-GPU session: warped-cell kernels -- parity tests, the variant comparison at config-3 size, ncu --set full of both kernels
set -u
mkdir -p gpurun_out
T=${TAG:-r02p}
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py tests/test_gpu_direct.py tests/test_zz_no_alloc.py -m gpu -q --tb=short -k "warped or wing or pair or no_alloc" > gpurun_out/${T}_pytest_warped.log 2>&1; echo "pytest warped rc=$?"
grep -E "passed|failed|Error|error" gpurun_out/${T}_pytest_warped.log | cut -c1-250 | tail -n 15
timeout 900 python scripts/warped_bench.py 816 > gpurun_out/${T}_warped_bench.json 2> gpurun_out/${T}_warped_bench.err; echo "warped bench rc=$?"
cut -c1-300 gpurun_out/${T}_warped_bench.err | tail -n 8
if [ "${NCU:-0}" = "1" ]; then
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_warped -s 6 -c 2 -f -o gpurun_out/${T}_warped python scripts/warped_bench.py 816 default > gpurun_out/${T}_ncu_warped.log 2>&1; echo "ncu warped rc=$?"
  ls -la gpurun_out/${T}_warped.ncu-rep
fi
