#!/bin/bash
# round 2, GPU call I (8 GPUs, tight timeouts): bench at N=8 and N=4 (whole-step graph with the NCCL all_reduces inside)
mkdir -p gpurun_out
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 \
    bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r2i_bench_n8.json 2> gpurun_out/r2i_bench_n8.err
echo "n8 rc=$?" >> gpurun_out/r2i_bench_n8.err
CUDA_VISIBLE_DEVICES=0,1,2,3 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29522 \
    bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r2i_bench_n4.json 2> gpurun_out/r2i_bench_n4.err
echo "n4 rc=$?" >> gpurun_out/r2i_bench_n4.err
tail -n 2 gpurun_out/r2i_bench_n8.err gpurun_out/r2i_bench_n4.err
