#!/bin/bash
# r2z: parallel fixed-order split-K reduction: conv / determinism / bench-config tests, per-call table
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_gpu.py tests/test_determinism_gpu.py tests/test_bench_config_gpu.py tests/test_head_gpu.py -m gpu -q --timeout 600 -x > gpurun_out/r2z_pytest.txt 2>&1
echo "pytest rc=$?"; tail -n 6 gpurun_out/r2z_pytest.txt
timeout 300 python scripts/call_table.py 64 > gpurun_out/r2z_call_table.txt 2>&1
head -n 12 gpurun_out/r2z_call_table.txt
timeout 300 python scripts/layer_table.py 64 > gpurun_out/r2z_layer.txt 2>&1
grep "wgrad" gpurun_out/r2z_layer.txt | head -n 24
