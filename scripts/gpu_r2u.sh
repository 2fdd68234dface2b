#!/bin/bash
# r2u: where the direct epilogue of the 16-channel layers spends its time (diagnostic build, knock-outs)
mkdir -p gpurun_out
export PDAC_B200_LIB=$PWD/pdac_pathological_image_segmentation_b200/libpdac_b200_prof.so
for d in 1 0; do
  for k in 0 1 2 4 8; do
    echo "== direct=$d HALO_DBG=$k"
    PDAC_HALO_DIRECT=$d HALO_DBG=$k timeout 120 python scripts/halo_prof.py 16 16 512 64 2>&1 | grep "conv3x3"
  done
done > gpurun_out/r2u_direct_knockouts.txt 2>&1
cat gpurun_out/r2u_direct_knockouts.txt
