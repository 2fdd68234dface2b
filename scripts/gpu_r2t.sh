#!/bin/bash
# r2t: direct (register -> global) epilogue of the N <= 32 halo kernels: full GPU suite, A/B layer table
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 -x > gpurun_out/r2t_pytest.txt 2>&1
echo "pytest rc=$?"; tail -n 15 gpurun_out/r2t_pytest.txt
timeout 300 python scripts/layer_table.py 64 > gpurun_out/r2t_layer_direct.txt 2>&1
PDAC_HALO_DIRECT=0 timeout 300 python scripts/layer_table.py 64 > gpurun_out/r2t_layer_staged.txt 2>&1
for f in direct staged; do echo "== $f"; grep "step\|512x512\|256x256 32->32" gpurun_out/r2t_layer_$f.txt; done
