#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests/test_head_gpu.py tests/test_smp_gpu.py tests/test_determinism_gpu.py tests/test_bench_config_gpu.py tests/test_wsi_gpu.py -m gpu -q --timeout 900 > gpurun_out/r2m_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2m_pytest.txt
python scripts/call_table.py 64 > gpurun_out/r2m_call_table.txt 2>&1
PDAC_HEAD_SCALAR=1 python scripts/call_table.py 64 > gpurun_out/r2m_call_table_scalar.txt 2>&1
python bench.py --steps 20 --warmup 3 --no-other-configs --no-cpu-baseline > gpurun_out/r2m_bench.json 2> gpurun_out/r2m_bench.err
echo "bench rc=$?" >> gpurun_out/r2m_bench.err
tail -n 12 gpurun_out/r2m_pytest.txt; tail -n 2 gpurun_out/r2m_bench.err; grep head gpurun_out/r2m_call_table.txt gpurun_out/r2m_call_table_scalar.txt; head -1 gpurun_out/r2m_call_table.txt
