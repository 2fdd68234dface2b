#!/bin/bash
# r3a: tap stacking on the N = 16 direct-epilogue layers: tests, standalone A/B, step table, WSI A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 -x > gpurun_out/r3a_pytest.txt 2>&1
echo "pytest rc=$?"; tail -n 8 gpurun_out/r3a_pytest.txt
for s in 1 0; do
  echo "== PDAC_HALO_STACK=$s"
  PDAC_HALO_STACK=$s PDAC_HALO_DEBUG=1 timeout 120 python scripts/one_conv.py 16 16 512 64 10 2>&1 | sort | uniq | tail -n 2
  PDAC_HALO_STACK=$s PDAC_HALO_DEBUG=1 timeout 120 python scripts/one_conv.py 32 16 512 64 10 2>&1 | sort | uniq | tail -n 2
done
timeout 300 python scripts/layer_table.py 64 > gpurun_out/r3a_layer_stack.txt 2>&1
PDAC_HALO_STACK=0 timeout 300 python scripts/layer_table.py 64 > gpurun_out/r3a_layer_nostack.txt 2>&1
for f in stack nostack; do echo "== $f"; grep "step\|512x512" gpurun_out/r3a_layer_$f.txt | grep -v wgrad; done
timeout 600 python bench.py --workload wsi --steps 5 --warmup 3 --no-cpu-baseline --no-other-configs > gpurun_out/r3a_wsi_stack.json 2> gpurun_out/r3a_wsi_stack.err
PDAC_HALO_STACK=0 timeout 600 python bench.py --workload wsi --steps 5 --warmup 3 --no-cpu-baseline --no-other-configs > gpurun_out/r3a_wsi_nostack.json 2> gpurun_out/r3a_wsi_nostack.err
for f in stack nostack; do python - <<PY
import json
try:
    d=json.loads(open("gpurun_out/r3a_wsi_$f.json").read().strip().splitlines()[-1]); print("$f", d["value"], d["unit"], d["ms_per_step"], d.get("e2e"))
except Exception as e: print("$f failed", e)
PY
done
