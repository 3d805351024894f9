#!/bin/bash
# 8 GPUs: the bench with its strong-scaling fit (SCALE's N = 8 point), NCCL topology logged
mkdir -p gpurun_out
NCCL_DEBUG=INFO NCCL_DEBUG_SUBSYS=INIT timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29534 \
    bench.py --gpus 8 --steps 10 --warmup 3 --no-gram --no-conv > gpurun_out/n8_bench.json 2> gpurun_out/n8_bench.err; echo "bench rc=$?" >> gpurun_out/n8_bench.err
grep -i "nvls\|Connected all\|Using network\|comm 0x" gpurun_out/n8_bench.err | head -20 > gpurun_out/n8_nccl.txt
