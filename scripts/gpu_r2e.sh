#!/bin/bash
# round 2, GPU call E (2 GPUs): NCCL tests, DDP timeline (overlap vs not), bench at N=2 with and without overlap
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/r2e_gpus.txt 2>&1
python -m pytest tests/test_trainer_gpu.py -m gpu -q --timeout 900 > gpurun_out/r2e_pytest.txt 2>&1
echo "pytest rc=$?" >> gpurun_out/r2e_pytest.txt
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
    scripts/ddp_timeline.py 64 > gpurun_out/r2e_timeline.txt 2>&1
echo "timeline rc=$?" >> gpurun_out/r2e_timeline.txt
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 \
    bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2e_bench_n2.json 2> gpurun_out/r2e_bench_n2.err
echo "bench rc=$?" >> gpurun_out/r2e_bench_n2.err
PDAC_DDP_OVERLAP=0 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 \
    bench.py --gpus 2 --steps 20 --warmup 3 > gpurun_out/r2e_bench_n2_nooverlap.json 2>> gpurun_out/r2e_bench_n2.err
python bench.py --steps 20 --warmup 3 --no-cpu-baseline --no-other-configs > gpurun_out/r2e_bench_n1.json 2>> gpurun_out/r2e_bench_n2.err
tail -n 4 gpurun_out/r2e_pytest.txt gpurun_out/r2e_timeline.txt gpurun_out/r2e_bench_n2.err
