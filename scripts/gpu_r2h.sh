#!/bin/bash
mkdir -p gpurun_out
export PDAC_B200_LIB=$PWD/pdac_pathological_image_segmentation_b200/libpdac_b200_prof.so
python scripts/halo_prof.py 64 64 128 64  128 128 64 64  16 16 512 64  256 256 32 64  512 512 16 64  32 16 512 64 > gpurun_out/r2h_halo_prof.txt 2>&1
HALO_DBG=1 python scripts/halo_prof.py 64 64 128 64 16 16 512 64 > gpurun_out/r2h_halo_prof_nostore.txt 2>&1
HALO_DBG=2 python scripts/halo_prof.py 64 64 128 64 16 16 512 64 > gpurun_out/r2h_halo_prof_epi_release_only.txt 2>&1
HALO_DBG=4 python scripts/halo_prof.py 64 64 128 64 16 16 512 64 > gpurun_out/r2h_halo_prof_no_mma.txt 2>&1
HALO_DBG=8 python scripts/halo_prof.py 64 64 128 64 16 16 512 64 > gpurun_out/r2h_halo_prof_no_stats.txt 2>&1
PDAC_HALO_DEBUG=1 python scripts/halo_prof.py 64 64 128 64  128 128 64 64  16 16 512 64 2>&1 | grep conv_halo | sort | uniq > gpurun_out/r2h_plans.txt
cat gpurun_out/r2h_halo_prof.txt
