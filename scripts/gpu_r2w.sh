#!/bin/bash
# r2w: direct epilogue, three non-spilling variants: full GPU suite, A/B per-layer table, standalone 16 -> 16
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 -x > gpurun_out/r2w_pytest.txt 2>&1
echo "pytest rc=$?"; tail -n 8 gpurun_out/r2w_pytest.txt
timeout 120 python scripts/one_conv.py 16 16 512 64 10
PDAC_HALO_DIRECT=0 timeout 120 python scripts/one_conv.py 16 16 512 64 10
timeout 120 python scripts/one_conv.py 32 16 512 64 10
PDAC_HALO_DIRECT=0 timeout 120 python scripts/one_conv.py 32 16 512 64 10
timeout 300 python scripts/layer_table.py 64 > gpurun_out/r2w_layer_direct.txt 2>&1
PDAC_HALO_DIRECT=0 timeout 300 python scripts/layer_table.py 64 > gpurun_out/r2w_layer_staged.txt 2>&1
for f in direct staged; do echo "== $f"; grep "step\|512x512\|256x256 32->32" gpurun_out/r2w_layer_$f.txt; done
