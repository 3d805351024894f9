#!/bin/bash
# round-2 GPU call 15: resident matvec at the fit's own shard sizes (1M rows = N = 1, 125k rows = N = 8), base vs default
mkdir -p gpurun_out
for rep in 1 2; do
 for n in 1000000 125000 262144; do
  echo "n $n base" >> gpurun_out/r15_cres.jsonl
  XGPR_B200_CRES_BASE=1 timeout 200 python tools/prof_cres.py $n 16384 >> gpurun_out/r15_cres.jsonl 2>&1
  echo "n $n default" >> gpurun_out/r15_cres.jsonl
  timeout 200 python tools/prof_cres.py $n 16384 >> gpurun_out/r15_cres.jsonl 2>&1
 done
done
