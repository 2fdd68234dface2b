#!/bin/bash
# r3c: direct pooled epilogue (mode 6) for the N <= 32 pooled data gradient: tests + A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_conv_gpu.py tests/test_determinism_gpu.py tests/test_smp_gpu.py tests/test_bench_config_gpu.py -m gpu -q --timeout 600 -x > gpurun_out/r3c_pytest.txt 2>&1
echo "pytest rc=$?"; tail -n 6 gpurun_out/r3c_pytest.txt
PDAC_HALO_DEBUG=1 timeout 300 python scripts/layer_table.py 64 > gpurun_out/r3c_layer.txt 2>&1
grep "step\|dgrad+pool" gpurun_out/r3c_layer.txt | grep -v "^\[conv" | head
grep "^\[conv_halo\] 64x512x512 K=1x16 N=32" gpurun_out/r3c_layer.txt | sort | uniq -c
