#!/bin/bash
# r2x: direct epilogue for every statistics-free halo launch (data gradients, inference): suite, A/B tables, WSI bench A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 -x > gpurun_out/r2x_pytest.txt 2>&1
echo "pytest rc=$?"; tail -n 8 gpurun_out/r2x_pytest.txt
timeout 300 python scripts/layer_table.py 64 > gpurun_out/r2x_layer_direct.txt 2>&1
PDAC_HALO_DIRECT=0 timeout 300 python scripts/layer_table.py 64 > gpurun_out/r2x_layer_staged.txt 2>&1
for f in direct staged; do echo "== $f"; grep "step\|dgrad" gpurun_out/r2x_layer_$f.txt | head -n 14; done
timeout 600 python bench.py --workload wsi --steps 5 --warmup 3 --no-cpu-baseline --no-other-configs > gpurun_out/r2x_wsi_direct.json 2> gpurun_out/r2x_wsi_direct.err
PDAC_HALO_DIRECT=0 timeout 600 python bench.py --workload wsi --steps 5 --warmup 3 --no-cpu-baseline --no-other-configs > gpurun_out/r2x_wsi_staged.json 2> gpurun_out/r2x_wsi_staged.err
for f in direct staged; do python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r2x_wsi_$f.json").read().strip().splitlines()[-1]); print("$f", d["value"], d["unit"], d["ms_per_step"], d.get("e2e"))
except Exception as e: print("$f failed", e)
PY
done
