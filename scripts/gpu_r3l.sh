#!/bin/bash
# r3l: inference: cached eval-mode BatchNorm affine, conv + BN + nearest x2 in one kernel: tests, WSI call table and bench
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r3l_pytest.txt 2>&1
echo "pytest rc=$?"; tail -n 6 gpurun_out/r3l_pytest.txt
timeout 300 python scripts/call_table_wsi.py 64 2>&1 | head -n 14
timeout 600 python bench.py --workload wsi --steps 5 --warmup 3 --no-cpu-baseline --no-other-configs > gpurun_out/r3l_wsi.json 2> gpurun_out/r3l_wsi.err
python - <<PY
import json
d=json.loads(open("gpurun_out/r3l_wsi.json").read().strip().splitlines()[-1]); print("wsi", d["value"], d["unit"], d["ms_per_step"], d.get("e2e"))
PY
