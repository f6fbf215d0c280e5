#!/bin/bash
set -u
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 15 gpurun_out/pytest_gpu.log
python bench.py --krylov-iters 20 > gpurun_out/bench.json 2> gpurun_out/bench.err; echo "bench rc=$?"; cat gpurun_out/bench.json; tail -n 5 gpurun_out/bench.err
FEMO_ASM_CHUNK=4 python bench.py --krylov-iters 5 --steps 5 > gpurun_out/bench_c4.json 2>&1; python -c "import json;d=json.loads(open('gpurun_out/bench_c4.json').read().strip().splitlines()[-1]);print('C=4',d['ms_per_step'],d['roofline']['avg_launch_ms'])"
FEMO_ASM_CHUNK=12 python bench.py --krylov-iters 5 --steps 5 > gpurun_out/bench_c12.json 2>&1; python -c "import json;d=json.loads(open('gpurun_out/bench_c12.json').read().strip().splitlines()[-1]);print('C=12',d['ms_per_step'],d['roofline']['avg_launch_ms'])"
FEMO_ASM_CHUNK=16 python bench.py --krylov-iters 5 --steps 5 > gpurun_out/bench_c16.json 2>&1; python -c "import json;d=json.loads(open('gpurun_out/bench_c16.json').read().strip().splitlines()[-1]);print('C=16',d['ms_per_step'],d['roofline']['avg_launch_ms'])"
timeout 600 python scripts/pcg_study.py 50 100 200 400 > gpurun_out/pcg_study.log 2>&1; cat gpurun_out/pcg_study.log
SHORT="python bench.py --steps 2 --warmup 3 --krylov-iters 3"
$SHORT > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:k_asm_planar -s 3 -c 1 -o gpurun_out/prof_asm $SHORT > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
