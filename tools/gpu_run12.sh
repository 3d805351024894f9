#!/bin/bash
# round-2 GPU call 12: resident matvec A/B -- default vs mixed widening vs mixed widening + L2 bulk prefetch
mkdir -p gpurun_out
for rep in 1 2; do
 for m in 16384 8192 4096; do
  timeout 200 python tools/prof_cres.py 262144 $m >> gpurun_out/r12_cres_default.jsonl 2>&1
  XGPR_B200_CRES_MIX=1 timeout 200 python tools/prof_cres.py 262144 $m >> gpurun_out/r12_cres_mix.jsonl 2>&1
  XGPR_B200_CRES_PF=1 timeout 200 python tools/prof_cres.py 262144 $m >> gpurun_out/r12_cres_pf.jsonl 2>&1
 done
done
XGPR_B200_CRES_PF=1 timeout 900 python -m pytest tests/test_stream_gpu.py tests/test_fitting_gpu.py -m gpu -q --timeout 600 > gpurun_out/r12_pytest_pf.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r12_pytest_pf.log
XGPR_B200_CRES_PF=1 timeout 300 python tools/prof_cres.py 262144 16384 > gpurun_out/plain_cres.log 2>&1 && \
XGPR_B200_CRES_PF=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:cres_zzv -s 3 -c 1 -f -o gpurun_out/prof12_cres python tools/prof_cres.py 262144 16384 > gpurun_out/ncu_cres.log 2>&1
