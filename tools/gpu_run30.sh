#!/bin/bash
# round-2 GPU call 30: ncu --set full of srht_sample_kernel<double> at the preconditioner's chunk shape
mkdir -p gpurun_out
timeout 200 python tools/prof_srht_sample.py > gpurun_out/r30_plain.json 2>&1 || exit 1
cat gpurun_out/r30_plain.json
timeout 900 ncu --set full --clock-control none --import-source on -k regex:srht_sample -s 3 -c 1 -f -o gpurun_out/prof30_srht_sample python tools/prof_srht_sample.py > gpurun_out/ncu_srht.log 2>&1
tail -2 gpurun_out/ncu_srht.log
