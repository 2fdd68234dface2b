#!/bin/bash
# r3o (8 GPUs): the N = 8 bench line with the final kernels
mkdir -p gpurun_out
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523 \
  bench.py --gpus 8 --steps 10 --warmup 3 --no-cpu-baseline --no-other-configs > gpurun_out/r3o_bench_n8.json 2> gpurun_out/r3o_bench_n8.err
echo "bench rc=$?"
python - <<PY
import json
d=json.loads(open("gpurun_out/r3o_bench_n8.json").read().strip().splitlines()[-1])
print(d["n_gpus"], d["value"], d["ms_per_step"], d["e2e"]["value"], d["clocks"])
PY
