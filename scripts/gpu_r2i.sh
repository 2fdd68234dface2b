#!/bin/bash
# round 2, GPU call I (8 GPUs, tight timeouts): train bench at N=8 / 4 / 2 and WSI bench at N=8 (whole-step graph with NCCL inside)
mkdir -p gpurun_out
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 \
    bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/r2i_bench_n8.json 2> gpurun_out/r2i_bench_n8.err
echo "n8 rc=$?" >> gpurun_out/r2i_bench_n8.err
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29523 \
    bench.py --gpus 8 --workload wsi --steps 20 --warmup 3 > gpurun_out/r2i_bench_wsi_n8.json 2>> gpurun_out/r2i_bench_n8.err
echo "wsi8 rc=$?" >> gpurun_out/r2i_bench_n8.err
CUDA_VISIBLE_DEVICES=0,1,2,3 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29522 \
    bench.py --gpus 4 --steps 20 --warmup 3 > gpurun_out/r2i_bench_n4.json 2> gpurun_out/r2i_bench_n4.err
echo "n4 rc=$?" >> gpurun_out/r2i_bench_n4.err
CUDA_VISIBLE_DEVICES=0,1 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29524 \
    bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2i_bench_n2.json 2> gpurun_out/r2i_bench_n2.err
echo "n2 rc=$?" >> gpurun_out/r2i_bench_n2.err
CUDA_VISIBLE_DEVICES=0 timeout 240 python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-other-configs > gpurun_out/r2i_bench_n1.json 2> gpurun_out/r2i_bench_n1.err
tail -n 1 gpurun_out/r2i_bench_n8.err gpurun_out/r2i_bench_n4.err gpurun_out/r2i_bench_n2.err
