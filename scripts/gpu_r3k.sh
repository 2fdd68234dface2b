#!/bin/bash
# r3k (2 GPUs): the 2-GPU NCCL test and the N = 2 bench line with the final kernels
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_trainer_gpu.py -m gpu -q --timeout 240 -k "two_gpu" > gpurun_out/r3k_pytest2.txt 2>&1
echo "pytest rc=$?"; tail -n 3 gpurun_out/r3k_pytest2.txt
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
  bench.py --gpus 2 --steps 10 --warmup 3 --no-cpu-baseline --no-other-configs > gpurun_out/r3k_bench_n2.json 2> gpurun_out/r3k_bench_n2.err
echo "bench rc=$?"; tail -c 900 gpurun_out/r3k_bench_n2.json | head -c 900; echo
