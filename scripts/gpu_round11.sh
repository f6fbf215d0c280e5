#!/bin/bash
# full round: all GPU tests, bench (ours + reference arm), smoke, ncu launch list, ncu full of the assembly kernel
set -u
mkdir -p gpurun_out
T=${TAG:-r01f}
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 4 gpurun_out/${T}_pytest_gpu.log
timeout 900 python bench.py --krylov-iters 50 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"; cat gpurun_out/${T}_bench.json; tail -n 3 gpurun_out/${T}_bench.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 1 > gpurun_out/${T}_bench_reference.json 2>&1; echo "ref rc=$?"; cat gpurun_out/${T}_bench_reference.json
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/${T}_smoke.log
SHORT="python bench.py --steps 2 --warmup 3 --krylov-iters 3 --no-mg"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_bench_steps2.csv $SHORT > gpurun_out/${T}_ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_asm_tiles -s 3 -c 1 -o gpurun_out/${T}_prof_asm $SHORT > gpurun_out/${T}_ncu_full.log 2>&1; echo "ncu full rc=$?"
ls -la gpurun_out | tail -n 12
