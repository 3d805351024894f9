#!/bin/bash
# 2 GPUs: NCCL-path parity test (sharded fit == single-GPU fit on the same rows) + the bench with its strong-scaling fit
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fitting_nccl_gpu.py -q --timeout 800 > gpurun_out/n2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/n2_pytest.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 \
    bench.py --gpus 2 --steps 10 --warmup 3 --no-gram --no-conv > gpurun_out/n2_bench.json 2> gpurun_out/n2_bench.err; echo "bench rc=$?" >> gpurun_out/n2_bench.err
