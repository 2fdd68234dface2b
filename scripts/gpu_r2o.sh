#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_head_gpu.py tests/test_determinism_gpu.py tests/test_smp_gpu.py tests/test_bench_config_gpu.py -m gpu -q --timeout 900 > gpurun_out/r2o_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2o_pytest.txt
python scripts/one_head.py 64 16 1 5 > gpurun_out/r2o_head_timing.txt 2>&1
python scripts/one_head.py 64 32 2 5 >> gpurun_out/r2o_head_timing.txt 2>&1
tail -n 4 gpurun_out/r2o_pytest.txt; cat gpurun_out/r2o_head_timing.txt
