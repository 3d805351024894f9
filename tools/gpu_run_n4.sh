#!/bin/bash
# 4 GPUs: the bench line as the driver's SCALE step launches it (N = 4 point)
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29535 \
    bench.py --gpus 4 --steps 10 --warmup 3 > gpurun_out/n4_bench.json 2> gpurun_out/n4_bench.err; echo "bench rc=$?" >> gpurun_out/n4_bench.err
tail -2 gpurun_out/n4_bench.err
grep "^{" gpurun_out/n4_bench.json | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value'], d['e2e']['value'], d['fit']['fit_s'], d['fit']['precond_s'], d['fit']['cg_s'], d['nmll']['value'], d['fit']['weights_vs_n1'])"
