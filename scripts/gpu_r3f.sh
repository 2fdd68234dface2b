#!/bin/bash
# r3f: tap stacking generalised (N <= 32, any chunk count, staged epilogue with BatchNorm sums = mode 7): tests + A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 -x > gpurun_out/r3f_pytest.txt 2>&1
echo "pytest rc=$?"; tail -n 6 gpurun_out/r3f_pytest.txt
for s in 1 0; do
  echo "== PDAC_HALO_STACK=$s"
  for shape in "128 32 256" "32 32 256" "64 32 256" "64 16 512"; do
    PDAC_HALO_STACK=$s PDAC_HALO_DEBUG=1 timeout 120 python scripts/one_conv.py $shape 64 10 2>&1 | sort | uniq | tail -n 2
  done
done
timeout 300 python scripts/layer_table.py 64 > gpurun_out/r3f_layer_stack.txt 2>&1
PDAC_HALO_STACK=0 timeout 300 python scripts/layer_table.py 64 > gpurun_out/r3f_layer_nostack.txt 2>&1
for f in stack nostack; do echo "== $f"; grep "step\|256x256\|512x512 32" gpurun_out/r3f_layer_$f.txt | grep -v wgrad; done
