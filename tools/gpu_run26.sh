#!/bin/bash
# round-2 GPU call 26: 16-byte Z^T t kernel behind both single-RHS consumers: tests, A/B, per-step times of a preconditioner chunk
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fitting_gpu.py tests/test_gpu_l1.py tests/test_scoring_gpu.py tests/test_stream_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r26_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/r26_tests.log
tail -3 gpurun_out/r26_tests.log
for v in 1 0; do
  if [ $v = 1 ]; then export XGPR_B200_ZTY_V1=1; else unset XGPR_B200_ZTY_V1; fi
  timeout 200 python tools/prof_zty.py 32768 16384 >> gpurun_out/r26_zty.jsonl 2>&1
  timeout 200 python tools/prof_zty.py 16384 32768 >> gpurun_out/r26_zty.jsonl 2>&1
done
unset XGPR_B200_ZTY_V1
cat gpurun_out/r26_zty.jsonl
timeout 300 python tools/prof_precond_chunk.py 32768 16384 >> gpurun_out/r26_precond_chunk.jsonl 2>&1
timeout 300 python tools/prof_precond_chunk.py 16384 32768 >> gpurun_out/r26_precond_chunk.jsonl 2>&1
cat gpurun_out/r26_precond_chunk.jsonl
