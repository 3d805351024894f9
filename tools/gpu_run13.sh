#!/bin/bash
# round-2 GPU call 13: resident matvec, L2 prefetch distance sweep
mkdir -p gpurun_out
for d in 1 2 3 4 6; do
 for m in 16384 8192 4096; do
  echo "dist $d" >> gpurun_out/r13_cres_pf.jsonl
  XGPR_B200_CRES_PF=1 XGPR_B200_CRES_PF_DIST=$d timeout 200 python tools/prof_cres.py 262144 $m >> gpurun_out/r13_cres_pf.jsonl 2>&1
 done
done
