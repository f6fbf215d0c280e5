#!/bin/bash
# after the SpMV rewrite: full GPU suite, bench, ncu full of k_spmv
set -u
mkdir -p gpurun_out
T=${TAG:-r01g}
timeout 1200 python -m pytest tests -m gpu -q > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 3 gpurun_out/${T}_pytest_gpu.log
timeout 900 python bench.py --krylov-iters 50 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"; cat gpurun_out/${T}_bench.json; tail -n 3 gpurun_out/${T}_bench.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/${T}_smoke.log
SHORT="python bench.py --steps 2 --warmup 3 --krylov-iters 3 --no-mg"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_bench_steps2.csv $SHORT > gpurun_out/${T}_ncu_list.log 2>&1; echo "ncu list rc=$?"
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_spmv -s 4 -c 1 -f -o gpurun_out/${T}_prof_spmv $SHORT > gpurun_out/${T}_ncu_full_spmv.log 2>&1; echo "ncu full rc=$?"
