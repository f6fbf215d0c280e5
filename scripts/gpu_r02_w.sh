#!/bin/bash
# N-GPU session: multi-GPU parity tests, then one strong-scaling bench line with the converged distributed multigrid solve
set -u
mkdir -p gpurun_out
T=${TAG:-r02w}
N=${NGPU:-4}
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -q --tb=short -rA -s > gpurun_out/${T}_pytest_dist.log 2>&1; echo "pytest dist rc=$?"
grep -E "passed|failed|^FAILED|^PASSED|^SKIPPED" gpurun_out/${T}_pytest_dist.log | cut -c1-250 | tail -n 30
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N"
for mode in ${MODES:-strong}; do
  FEMO_VERBOSE=1 FEMO_TRACE_OVERLAP=1 timeout 600 $RUN --scaling $mode --steps 5 --warmup 3 --krylov-iters 300 > gpurun_out/${T}_bench_${N}gpu_${mode}.json 2> gpurun_out/${T}_bench_${N}gpu_${mode}.err
  echo "bench $mode rc=$?"
done
python - <<PY
import json
for m in "${MODES:-strong}".split():
    try:
        d=json.loads(open("gpurun_out/${T}_bench_${N}gpu_%s.json"%m).read().strip().splitlines()[-1])
        mg=d["solve"]["multigrid"] or {}
        print(m, "asm ms %.3f"%d["ms_per_step"], "GDOF/s %.2f"%d["value"], "e2e %.2f"%d["e2e"]["value"], "pcg ms/iter %.4f"%(1e3*d["solve"]["s_per_krylov_iter"]),
              "overlap", d["solve"].get("halo_overlap_trace"))
        print("   mg", {k:(round(v,4) if isinstance(v,float) else v) for k,v in mg.items() if k in ("state_iters","state_s","adjoint_iters","adjoint_s","setup_s","ms_per_vcycle_iteration","n_dof","n_dist","error","hierarchy_build_s")})
    except Exception as e:
        print(m, "unreadable", e)
PY
