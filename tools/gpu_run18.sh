#!/bin/bash
# round-2 GPU call 18: full GPU suite with the encoded multi-RHS path, then the default bench (now with the nmll record)
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r18_pytest_gpu.log 2>&1
echo "suite exit $?" >> gpurun_out/r18_pytest_gpu.log
tail -5 gpurun_out/r18_pytest_gpu.log
( time timeout 900 python bench.py > gpurun_out/r18_bench_default.json 2> gpurun_out/r18_bench_default.err ) 2> gpurun_out/r18_bench_time.txt
tail -c 1500 gpurun_out/r18_bench_default.json; cat gpurun_out/r18_bench_time.txt; tail -5 gpurun_out/r18_bench_default.err
