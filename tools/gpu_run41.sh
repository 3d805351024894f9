#!/bin/bash
# round-2 GPU call 41: verification of the final tree -- full GPU suite, smoke, default bench, reference arm
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r41_pytest_gpu.log 2>&1
echo "suite exit $?" >> gpurun_out/r41_pytest_gpu.log
tail -4 gpurun_out/r41_pytest_gpu.log
timeout 600 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/r41_smoke.log 2>&1
tail -2 gpurun_out/r41_smoke.log
timeout 900 python bench.py > gpurun_out/r41_bench_default.json 2> gpurun_out/r41_bench_default.err
echo "bench exit $?"
timeout 900 python bench.py --impl reference > gpurun_out/r41_bench_reference.json 2> gpurun_out/r41_bench_reference.err
echo "reference exit $?"
tail -c 600 gpurun_out/r41_bench_reference.json
