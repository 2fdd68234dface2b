#!/bin/bash
# r2p: the allocation-free entry point alone in a fresh process, then after other tests; the loss-curve parity test.
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -m gpu -q --timeout 200 > gpurun_out/r2p_zz_alone.txt 2>&1
echo "zz alone rc=$?"; tail -n 12 gpurun_out/r2p_zz_alone.txt
timeout 600 python -m pytest tests/test_elementwise_gpu.py tests/test_determinism_gpu.py tests/test_zz_no_alloc.py -m gpu -q --timeout 300 > gpurun_out/r2p_zz_after.txt 2>&1
echo "zz after others rc=$?"; tail -n 12 gpurun_out/r2p_zz_after.txt
timeout 600 python -m pytest tests/test_trainer_gpu.py -m gpu -q --timeout 300 -k "loss_curve or validate" > gpurun_out/r2p_trainer.txt 2>&1
echo "trainer rc=$?"; tail -n 25 gpurun_out/r2p_trainer.txt
