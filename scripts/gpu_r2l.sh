#!/bin/bash
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r2l_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2l_pytest.txt
python scripts/call_table.py 64 > gpurun_out/r2l_call_table.txt 2>&1
python bench.py --steps 20 --warmup 3 --no-other-configs > gpurun_out/r2l_bench.json 2> gpurun_out/r2l_bench.err
echo "bench rc=$?" >> gpurun_out/r2l_bench.err
python bench.py --workload wsi --steps 20 --warmup 3 > gpurun_out/r2l_bench_wsi.json 2>> gpurun_out/r2l_bench.err
tail -n 12 gpurun_out/r2l_pytest.txt; tail -n 2 gpurun_out/r2l_bench.err; head -30 gpurun_out/r2l_call_table.txt
