#!/bin/bash
# r3d: where the pooled 16 -> 32 data gradient spends its time (diagnostic build, knock-outs), direct vs staged epilogue
mkdir -p gpurun_out
export PDAC_B200_LIB=$PWD/pdac_pathological_image_segmentation_b200/libpdac_b200_prof.so
for d in 1 0; do for k in 0 2 4 1; do
  echo "== direct=$d HALO_DBG=$k"
  HALO_OP=pool PDAC_HALO_DIRECT=$d HALO_DBG=$k timeout 120 python scripts/halo_prof.py 16 32 512 64 2>&1 | grep "conv3x3"
done; done > gpurun_out/r3d_pool_knockouts.txt 2>&1
cat gpurun_out/r3d_pool_knockouts.txt
