#!/bin/bash
# round 2, GPU call G: full GPU suite, row-shift level 1 vs 2, default bench (all records), WSI bench
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 900 > gpurun_out/r2g_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2g_pytest.txt
python scripts/layer_table.py 64 > gpurun_out/r2g_layer_table.txt 2>&1
PDAC_WGRAD_ROWSHIFT=1 python scripts/layer_table.py 64 > gpurun_out/r2g_layer_table_rs1.txt 2>&1
python scripts/call_table.py 64 > gpurun_out/r2g_call_table.txt 2>&1
python bench.py --steps 20 --warmup 3 > gpurun_out/r2g_bench.json 2> gpurun_out/r2g_bench.err
echo "bench rc=$?" >> gpurun_out/r2g_bench.err
python bench.py --workload wsi --steps 20 --warmup 3 > gpurun_out/r2g_bench_wsi.json 2>> gpurun_out/r2g_bench.err
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2g_bench_reference.json 2>> gpurun_out/r2g_bench.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2g_smoke.txt 2>&1
tail -n 3 gpurun_out/r2g_pytest.txt gpurun_out/r2g_bench.err gpurun_out/r2g_smoke.txt
