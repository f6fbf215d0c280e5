#!/bin/bash
# 2-GPU session of round 2: multi-GPU parity tests, then weak / strong bench lines at small and full size.
set -u
mkdir -p gpurun_out
T=${TAG:-r02d}
N=${NGPU:-2}
timeout 900 python -m pytest tests/test_gpu_dist.py -m gpu -q --tb=short -rA -s > gpurun_out/${T}_pytest_dist.log 2>&1; echo "pytest dist rc=$?"
grep -E "^rank|passed|failed|^FAILED|^PASSED" gpurun_out/${T}_pytest_dist.log | cut -c1-330 | tail -n 40
RUN="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29544 bench.py --gpus $N"
for mode in weak strong; do
  FEMO_VERBOSE=1 timeout 900 $RUN --scaling $mode --steps 5 --warmup 3 --krylov-iters 200 > gpurun_out/${T}_bench_${N}gpu_${mode}.json 2> gpurun_out/${T}_bench_${N}gpu_${mode}.err
  echo "bench $mode rc=$?"; cat gpurun_out/${T}_bench_${N}gpu_${mode}.json; grep -v "tile plan\|OMP_NUM\|\*\*\*\*" gpurun_out/${T}_bench_${N}gpu_${mode}.err | tail -n 8
done
# the classic two-reduction, non-overlapped loop for comparison (strong scaling is where it matters)
FEMO_PCG_CLASSIC=1 FEMO_HALO_OVERLAP=0 timeout 600 $RUN --scaling strong --steps 3 --warmup 3 --krylov-iters 200 --no-mg > gpurun_out/${T}_bench_${N}gpu_strong_classic.json 2> gpurun_out/${T}_bench_${N}gpu_strong_classic.err
echo "bench strong classic rc=$?"; python - <<PY
import json
for m in ("weak","strong","strong_classic"):
    try:
        d=json.loads(open("gpurun_out/${T}_bench_${N}gpu_%s.json"%m).read().strip().splitlines()[-1])
        print(m, "pcg ms/iter", 1e3*d["solve"]["s_per_krylov_iter"], "asm ms", d["ms_per_step"])
    except Exception as e:
        print(m, "unreadable", e)
PY
