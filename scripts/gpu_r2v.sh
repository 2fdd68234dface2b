#!/bin/bash
# r2v: ncu source-level profile of the direct epilogue on the 16 -> 16 layer (after the same command ran clean in r2u)
mkdir -p gpurun_out
timeout 120 python scripts/one_conv.py 16 16 512 64 > gpurun_out/r2v_plain.txt 2>&1 || exit 1
cat gpurun_out/r2v_plain.txt
timeout 600 ncu --set full --clock-control none --import-source on -k regex:"conv_halo" -s 2 -c 1 -o /tmp/r2v python scripts/one_conv.py 16 16 512 64 1 > gpurun_out/r2v_ncu.log 2>&1
ncu -i /tmp/r2v.ncu-rep --page details > gpurun_out/r2v_details.txt 2>/dev/null
ncu -i /tmp/r2v.ncu-rep --page source --csv > gpurun_out/r2v_source.csv 2>/dev/null
ls -la gpurun_out/r2v_*; tail -n 3 gpurun_out/r2v_ncu.log
