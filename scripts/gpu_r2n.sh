#!/bin/bash
mkdir -p gpurun_out
python scripts/one_head.py 64 16 1 1 > gpurun_out/r2n_head_timing.txt 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:"head3x3_fwd" -s 1 -c 1 -o /tmp/r2n_head python scripts/one_head.py 64 16 1 1 > gpurun_out/r2n_ncu.log 2>&1
ncu -i /tmp/r2n_head.ncu-rep --page details > gpurun_out/r2n_fwd_details.txt 2>/dev/null
ncu -i /tmp/r2n_head.ncu-rep --page source --csv > gpurun_out/r2n_fwd_source.csv 2>/dev/null
cat gpurun_out/r2n_head_timing.txt
