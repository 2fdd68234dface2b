#!/bin/bash
# round 2, GPU call C: full GPU test suite, bench, ncu launch list and --set full of the three top kernels at B=64
mkdir -p gpurun_out
python -m pytest tests -m gpu -q --timeout 1800 > gpurun_out/r2c_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2c_pytest.txt
python bench.py --steps 10 --warmup 3 --no-other-configs > gpurun_out/r2c_bench.json 2> gpurun_out/r2c_bench.err
python scripts/call_table.py 64 pdac_bn_prepare > gpurun_out/r2c_call_table.txt 2>&1
python scripts/layer_table.py 64 > gpurun_out/r2c_layer_table.txt 2>&1
python scripts/one_step.py 64 3 > gpurun_out/r2c_plain.log 2>&1 &&
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
    --log-file gpurun_out/r2c_launches_raw.csv python scripts/one_step.py 64 3 > gpurun_out/r2c_ncu_launches.log 2>&1
python scripts/one_step.py 64 3 >> gpurun_out/r2c_plain.log 2>&1 &&
ncu --set full --clock-control none -k regex:"conv_halo_kernel|wgrad_halo_kernel|bn_bwd_apply_kernel" -s 214 -c 107 \
    -o /tmp/r2c_full python scripts/one_step.py 64 3 > gpurun_out/r2c_ncu_full.log 2>&1
ncu -i /tmp/r2c_full.ncu-rep --page raw --csv > gpurun_out/r2c_full_raw.csv 2>/dev/null
du -sh gpurun_out
tail -n 3 gpurun_out/r2c_pytest.txt
