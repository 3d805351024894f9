#!/bin/bash
# round-2 GPU call 22: ncu --set full of tc_gram_pair_kernel (cta_group::2) at 32,768 x 8,192, 5 x 8-bit planes
mkdir -p gpurun_out
XGPR_B200_TC_PAIR=1 timeout 200 python tools/prof_gram_tc.py > gpurun_out/r22_plain.txt 2>&1 || exit 1
cat gpurun_out/r22_plain.txt
XGPR_B200_TC_PAIR=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:tc_gram -s 2 -c 1 -f -o gpurun_out/prof22_gram_pair python tools/prof_gram_tc.py > gpurun_out/ncu_gram_pair.log 2>&1
tail -3 gpurun_out/ncu_gram_pair.log
