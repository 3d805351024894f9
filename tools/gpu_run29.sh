#!/bin/bash
# round-2 GPU call 29: A^T Z sketch GEMM with the last partial wave split along K (128 tiles on 148 SMs)
mkdir -p gpurun_out
for v in 0 1 0 1; do
  XGPR_B200_TC_SPLIT=$v timeout 300 python tools/prof_precond_chunk.py 32768 16384 >> gpurun_out/r29_split.jsonl 2>&1
done
XGPR_B200_TC_SPLIT=1 timeout 300 python tools/prof_precond_chunk.py 16384 32768 >> gpurun_out/r29_split.jsonl 2>&1
XGPR_B200_TC_SPLIT=0 timeout 300 python tools/prof_precond_chunk.py 16384 32768 >> gpurun_out/r29_split.jsonl 2>&1
cat gpurun_out/r29_split.jsonl
XGPR_B200_TC_SPLIT=1 timeout 600 python -m pytest tests/test_gram_tc_gpu.py tests/test_fitting_gpu.py -x -q -m gpu 2>&1 | tail -3
