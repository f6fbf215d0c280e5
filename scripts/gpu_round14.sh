#!/bin/bash
# round r01i: full GPU suite (static + dynamic + shape partials), bench both arms, smoke, ncu launch list, config-4 timings
set -u
mkdir -p gpurun_out
T=${TAG:-r01i}
timeout 1500 python -m pytest tests -m gpu -q --tb=short > gpurun_out/${T}_pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -n 40 gpurun_out/${T}_pytest_gpu.log
timeout 600 python scripts/dyn_bench.py --large --out gpurun_out/${T}_dyn_bench.json > gpurun_out/${T}_dyn_bench.log 2>&1; echo "dyn bench rc=$?"; tail -n 12 gpurun_out/${T}_dyn_bench.log
timeout 900 python bench.py --krylov-iters 50 > gpurun_out/${T}_bench.json 2> gpurun_out/${T}_bench.err; echo "bench rc=$?"; cat gpurun_out/${T}_bench.json; tail -n 3 gpurun_out/${T}_bench.err
timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/${T}_bench_reference.json 2>> gpurun_out/${T}_bench.err; echo "bench ref rc=$?"; cat gpurun_out/${T}_bench_reference.json
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/${T}_smoke.log 2>&1; echo "smoke rc=$?"; tail -n 2 gpurun_out/${T}_smoke.log
SHORT="python bench.py --steps 2 --warmup 3 --krylov-iters 3 --no-mg"
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/${T}_launches_bench_steps2.csv $SHORT > gpurun_out/${T}_ncu_list.log 2>&1; echo "ncu list rc=$?"
