#!/bin/bash
# r3n (4 GPUs): the N = 4 bench line with the final kernels
mkdir -p gpurun_out
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29519 \
  bench.py --gpus 4 --steps 10 --warmup 3 --no-cpu-baseline --no-other-configs > gpurun_out/r3n_bench_n4.json 2> gpurun_out/r3n_bench_n4.err
echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/r3n_bench_n4.json").read().strip().splitlines()[-1])
print(d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"], d["clocks"])
PY
