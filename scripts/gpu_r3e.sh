#!/bin/bash
# r3e: packed maxpool forward: tests + per-call table
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_elementwise_gpu.py tests/test_resunet_gpu.py tests/test_determinism_gpu.py tests/test_bench_config_gpu.py -m gpu -q --timeout 600 -x > gpurun_out/r3e_pytest.txt 2>&1
echo "pytest rc=$?"; tail -n 6 gpurun_out/r3e_pytest.txt
timeout 300 python scripts/call_table.py 64 > gpurun_out/r3e_call_table.txt 2>&1
grep "step\|maxpool\|conv_halo" gpurun_out/r3e_call_table.txt
