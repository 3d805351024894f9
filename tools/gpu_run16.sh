#!/bin/bash
# round-2 GPU call 16: the allocation-free matvec entry point (alone in a fresh process, then after the whole suite), and the
# resident matvec with batches dealt round-robin to the clusters (XGPR_B200_CRES_INTERLEAVE) at the fit's shard sizes
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r16_noalloc_alone.log 2>&1
echo "alone exit $?" >> gpurun_out/r16_noalloc_alone.log
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r16_pytest_gpu.log 2>&1
echo "suite exit $?" >> gpurun_out/r16_pytest_gpu.log
timeout 600 python -m pytest tests/test_stream_gpu.py tests/test_fitting_gpu.py -x -q -m gpu > gpurun_out/r16_interleave_tests.log 2>&1
echo "interleave exit $?" >> gpurun_out/r16_interleave_tests.log
for rep in 1 2; do
 for n in 1000000 262144 125000; do
  echo "n $n default" >> gpurun_out/r16_cres.jsonl
  timeout 200 python tools/prof_cres.py $n 16384 >> gpurun_out/r16_cres.jsonl 2>&1
  echo "n $n interleave(default)" >> gpurun_out/r16_cres.jsonl
  timeout 200 python tools/prof_cres.py $n 16384 >> gpurun_out/r16_cres.jsonl 2>&1
 done
done
for n in 500000 1000000; do
  echo "n $n m 8192 default" >> gpurun_out/r16_cres.jsonl
  timeout 200 python tools/prof_cres.py $n 8192 >> gpurun_out/r16_cres.jsonl 2>&1
  echo "n $n m 8192 interleave(default)" >> gpurun_out/r16_cres.jsonl
  timeout 200 python tools/prof_cres.py $n 8192 >> gpurun_out/r16_cres.jsonl 2>&1
done
tail -3 gpurun_out/r16_noalloc_alone.log gpurun_out/r16_pytest_gpu.log gpurun_out/r16_interleave_tests.log
cat gpurun_out/r16_cres.jsonl
