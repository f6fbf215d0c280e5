#!/bin/bash
# 2-GPU session: multi-GPU solve parity (incl. the one-call residual evaluation on a rank), then the driver-style weak bench line
set -u
mkdir -p gpurun_out
T=${TAG:-r02y2}
timeout 600 python -m pytest tests/test_gpu_dist.py -m gpu -q --tb=short -rA -s -k "two_gpu_solve" > gpurun_out/${T}_pytest_dist.log 2>&1; echo "pytest dist rc=$?"
grep -E "passed|failed|^FAILED|^PASSED" gpurun_out/${T}_pytest_dist.log | cut -c1-200 | tail -n 8
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/${T}_bench_2gpu_weak.json 2> gpurun_out/${T}_bench_2gpu_weak.err; echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/${T}_bench_2gpu_weak.json").read().strip().splitlines()[-1])
mg=d["solve"]["multigrid"]
print("value %.2f"%d["value"], "e2e %.3f (%.2f ms)"%(d["e2e"]["value"], d["e2e"]["ms_per_step"]), "separate %.3f (%.2f ms)"%(d["e2e"]["separate_calls"]["value"], d["e2e"]["separate_calls"]["ms_per_step"]),
      "| mg setup %.3f build %.1f state %.3f"%(mg["setup_s"], mg["hierarchy_build_s"], mg["state_s"]))
PY
