#!/bin/bash
# round 2, GPU call D: row-shift weight gradient — tests, layer table with / without it, bench; then compute-sanitizer memcheck
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 1800 > gpurun_out/r2d_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2d_pytest.txt
python scripts/layer_table.py 64 > gpurun_out/r2d_layer_table.txt 2>&1
PDAC_WGRAD_ROWSHIFT=0 python scripts/layer_table.py 64 > gpurun_out/r2d_layer_table_norowshift.txt 2>&1
python bench.py --steps 10 --warmup 3 --no-other-configs > gpurun_out/r2d_bench.json 2> gpurun_out/r2d_bench.err
timeout 900 compute-sanitizer --tool memcheck --error-exitcode 7 python -m pytest tests/test_conv_gpu.py tests/test_elementwise_gpu.py -m gpu -q -x --timeout 800 \
    -k "not 512" > gpurun_out/r2d_memcheck.txt 2>&1
echo "memcheck rc=$?" >> gpurun_out/r2d_memcheck.txt
tail -n 3 gpurun_out/r2d_pytest.txt gpurun_out/r2d_memcheck.txt
