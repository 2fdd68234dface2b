#!/bin/bash
# round 2, GPU call A: full GPU test suite, gradient-conditioning diagnostic, bench (deterministic and atomics), call tables
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm --format=csv > gpurun_out/r2a_gpu.txt 2>&1
python -m pytest tests -m gpu -q --timeout 1200 -x --deselect tests/test_bench_config_gpu.py > gpurun_out/r2a_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2a_pytest.txt
python -m pytest tests/test_bench_config_gpu.py -m gpu -q --timeout 1800 > gpurun_out/r2a_pytest_bench_config.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2a_pytest_bench_config.txt
python scripts/diag_gradcond.py 64 2 > gpurun_out/r2a_gradcond.txt 2>&1
python scripts/diag_gradcond.py 256 4 >> gpurun_out/r2a_gradcond.txt 2>&1
python bench.py --steps 10 --warmup 3 > gpurun_out/r2a_bench.json 2> gpurun_out/r2a_bench.err
echo "bench rc=$?" >> gpurun_out/r2a_bench.err
PDAC_NONDETERMINISTIC=1 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-other-configs > gpurun_out/r2a_bench_atomics.json 2>> gpurun_out/r2a_bench.err
python scripts/call_table.py 64 pdac_bn_bwd_reduce pdac_bn_prepare > gpurun_out/r2a_call_table.txt 2>&1
python scripts/layer_table.py 64 > gpurun_out/r2a_layer_table.txt 2>&1
PDAC_NONDETERMINISTIC=1 python scripts/layer_table.py 64 > gpurun_out/r2a_layer_table_atomics.txt 2>&1
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2a_smoke.txt 2>&1
tail -3 gpurun_out/r2a_pytest.txt gpurun_out/r2a_pytest_bench_config.txt gpurun_out/r2a_bench.err
