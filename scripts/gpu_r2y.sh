#!/bin/bash
# r2y: upper bound of tap stacking on the N = 16 layers: direct epilogue with the MMAs knocked out (HALO_DBG=4)
mkdir -p gpurun_out
export PDAC_B200_LIB=$PWD/pdac_pathological_image_segmentation_b200/libpdac_b200_prof.so
for k in 0 4 2; do
  echo "== HALO_DBG=$k"
  HALO_DBG=$k timeout 120 python scripts/halo_prof.py 16 16 512 64  32 16 512 64  32 32 256 64 2>&1 | grep "conv3x3"
done > gpurun_out/r2y_direct_knockouts.txt 2>&1
cat gpurun_out/r2y_direct_knockouts.txt
