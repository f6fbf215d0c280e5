#!/bin/bash
# N-GPU scaling session of round 2 (charged N x): weak and strong bench lines with the converged distributed multigrid solve and
# the CUDA-event trace of the halo / SpMV overlap.   TAG=r02m NGPU=8 bash scripts/gpu_r02_scale.sh
set -u
mkdir -p gpurun_out
T=${TAG:-r02m}
N=${NGPU:-8}
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N"
for mode in strong weak; do
  FEMO_VERBOSE=1 FEMO_TRACE_OVERLAP=1 timeout 600 $RUN --scaling $mode --steps 5 --warmup 3 --krylov-iters 300 > gpurun_out/${T}_bench_${N}gpu_${mode}.json 2> gpurun_out/${T}_bench_${N}gpu_${mode}.err
  echo "bench $mode rc=$?"
done
python - <<PY
import json
for m in ("strong","weak"):
    try:
        d=json.loads(open("gpurun_out/${T}_bench_${N}gpu_%s.json"%m).read().strip().splitlines()[-1])
        mg=d["solve"]["multigrid"] or {}
        print(m, "asm ms %.3f"%d["ms_per_step"], "GDOF/s %.2f"%d["value"], "e2e %.2f"%d["e2e"]["value"], "pcg ms/iter %.4f"%(1e3*d["solve"]["s_per_krylov_iter"]),
              "overlap", d["solve"].get("halo_overlap_trace"))
        print("   mg", {k:(round(v,4) if isinstance(v,float) else v) for k,v in mg.items() if k in ("state_iters","state_s","adjoint_iters","adjoint_s","setup_s","ms_per_vcycle_iteration","n_dof","error","hierarchy_build_s")})
    except Exception as e:
        print(m, "unreadable", e)
PY
grep -h "mg level 0\|mg setup" gpurun_out/${T}_bench_${N}gpu_strong.err | head -4
