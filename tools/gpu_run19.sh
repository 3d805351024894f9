#!/bin/bash
# round-2 GPU call 19: full GPU suite, cfg5 NMLL phases, default bench with the nmll record
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r19_pytest_gpu.log 2>&1
echo "suite exit $?" >> gpurun_out/r19_pytest_gpu.log
tail -5 gpurun_out/r19_pytest_gpu.log
timeout 300 python tools/bench_nmll.py --rows 2000000 > gpurun_out/r19_nmll.json 2>&1
tail -c 900 gpurun_out/r19_nmll.json
( time timeout 900 python bench.py > gpurun_out/r19_bench_default.json 2> gpurun_out/r19_bench_default.err ) 2> gpurun_out/r19_bench_time.txt
cat gpurun_out/r19_bench_time.txt; tail -5 gpurun_out/r19_bench_default.err
