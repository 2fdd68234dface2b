#!/bin/bash
# r2r: fused nearest-x2 backward in the data-gradient epilogue: kernel parity, step-level bitwise test, step time
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_conv_gpu.py tests/test_determinism_gpu.py -m gpu -q --timeout 300 -x -k "pool2 or fused_upsampling" > gpurun_out/r2r_pytest.txt 2>&1
echo "pytest rc=$?"; tail -n 30 gpurun_out/r2r_pytest.txt
PDAC_HALO_DEBUG=1 timeout 300 python scripts/call_table.py 64 pdac_upsample_nearest2x_bwd > gpurun_out/r2r_call_table.txt 2>&1
grep -v "^\[conv_halo\]" gpurun_out/r2r_call_table.txt | head -n 30
grep "^\[conv_halo\]" gpurun_out/r2r_call_table.txt | sort | uniq -c | sort -rn | head -n 60 > gpurun_out/r2r_plans.txt
