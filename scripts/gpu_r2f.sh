#!/bin/bash
# round 2, GPU call F (2 GPUs, tight timeouts): NCCL test, bench at N=2 with / without overlap
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_trainer_gpu.py -m gpu -q --timeout 240 -k two_gpu > gpurun_out/r2f_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2f_pytest.txt
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2f_bench_n2.json 2> gpurun_out/r2f_bench_n2.err
echo "bench rc=$?" >> gpurun_out/r2f_bench_n2.err
PDAC_DDP_OVERLAP=0 timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2f_bench_n2_nooverlap.json 2>> gpurun_out/r2f_bench_n2.err
echo "bench2 rc=$?" >> gpurun_out/r2f_bench_n2.err
timeout 240 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29514 \
    bench.py --gpus 2 --workload wsi --steps 10 --warmup 3 > gpurun_out/r2f_bench_wsi_n2.json 2>> gpurun_out/r2f_bench_n2.err
echo "wsi rc=$?" >> gpurun_out/r2f_bench_n2.err
tail -n 5 gpurun_out/r2f_pytest.txt gpurun_out/r2f_bench_n2.err
