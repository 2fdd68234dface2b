#!/bin/bash
# round 2, GPU call J: fused up2 in the normalise pass, double-buffered dual epilogue, small tail bucket — tests + tables + bench
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r2j_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2j_pytest.txt
python scripts/layer_table.py 64 > gpurun_out/r2j_layer_table.txt 2>&1
python scripts/call_table.py 64 > gpurun_out/r2j_call_table.txt 2>&1
python bench.py --steps 20 --warmup 3 --no-other-configs > gpurun_out/r2j_bench.json 2> gpurun_out/r2j_bench.err
echo "bench rc=$?" >> gpurun_out/r2j_bench.err
tail -n 3 gpurun_out/r2j_pytest.txt gpurun_out/r2j_bench.err
