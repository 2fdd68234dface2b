#!/bin/bash
# r3h: does tap stacking pay at N = 64 (one 192-column accumulator per stage, 128-row tiles)?
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 -x > gpurun_out/r3h_pytest.txt 2>&1
echo "pytest rc=$?"; tail -n 4 gpurun_out/r3h_pytest.txt
PDAC_HALO_STACK_MAXN=64 timeout 600 python -m pytest tests/test_conv_gpu.py tests/test_smp_gpu.py tests/test_resunet_gpu.py -m gpu -q --timeout 600 > gpurun_out/r3h_pytest64.txt 2>&1
echo "pytest (MAXN=64) rc=$?"; tail -n 4 gpurun_out/r3h_pytest64.txt
for n in 64 32; do
  echo "== PDAC_HALO_STACK_MAXN=$n"
  for shape in "64 64 128" "192 64 128" "128 64 128"; do
    PDAC_HALO_STACK_MAXN=$n PDAC_HALO_DEBUG=1 timeout 120 python scripts/one_conv.py $shape 64 10 2>&1 | sort | uniq | tail -n 2
  done
done
PDAC_HALO_STACK_MAXN=64 timeout 300 python scripts/layer_table.py 64 > gpurun_out/r3h_layer_64.txt 2>&1
timeout 300 python scripts/layer_table.py 64 > gpurun_out/r3h_layer_32.txt 2>&1
for f in 64 32; do echo "== MAXN=$f"; grep "step\|128x128" gpurun_out/r3h_layer_$f.txt | grep -v wgrad; done
