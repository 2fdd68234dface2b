#!/bin/bash
# r3j: final measurement refresh of round 2 (same recipe as r3b) + smoke()
# B=64 step and --set full of every conv_halo launch of that step (reports stay in /tmp, csv exported)
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r3j_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r3j_pytest.txt; tail -n 3 gpurun_out/r3j_pytest.txt
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/r3j_bench.json 2> gpurun_out/r3j_bench.err
tail -c 600 gpurun_out/r3j_bench.json; echo
timeout 300 python scripts/call_table.py 64 > gpurun_out/r3j_call_table.txt 2>&1
timeout 300 python scripts/layer_table.py 64 > gpurun_out/r3j_layer_table.txt 2>&1
timeout 300 python scripts/one_step.py 64 3 > gpurun_out/r3j_plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --clock-control none --csv \
    --log-file gpurun_out/r3j_launches_raw.csv python scripts/one_step.py 64 3 > gpurun_out/r3j_ncu_launches.log 2>&1
n=$(grep -c "conv_halo_kernel" gpurun_out/r3j_launches_raw.csv); per=$((n / 9)); echo "conv_halo launches per step: $per"
timeout 900 ncu --set full --clock-control none -k regex:"conv_halo_kernel" -s $((2 * per)) -c $per \
    -o /tmp/r3j_full python scripts/one_step.py 64 3 > gpurun_out/r3j_ncu_full.log 2>&1
ncu -i /tmp/r3j_full.ncu-rep --page raw --csv > gpurun_out/r3j_full_raw.csv 2>/dev/null
du -sh gpurun_out; tail -n 2 gpurun_out/r3j_ncu_full.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r3j_smoke.txt 2>&1; tail -n 2 gpurun_out/r3j_smoke.txt
