#!/bin/bash
# round-2 GPU call 20: ncu --set full of the encoded 24+2 multi-RHS kernels (zv_mma / ztt_mma <float, tail>)
mkdir -p gpurun_out
timeout 200 python tools/prof_multirhs_c32.py 16384 32768 26 > gpurun_out/r20_plain.json 2>&1 || exit 1
cat gpurun_out/r20_plain.json
# launches of the tool: 3 f64 pairs... skip to the c32 pairs: kernel-name filter on the float instantiations
timeout 900 ncu --set full --clock-control none --import-source on -k regex:mma_kernel -s 18 -c 2 -f -o gpurun_out/prof20_multirhs python tools/prof_multirhs_c32.py 16384 32768 26 > gpurun_out/ncu_multirhs.log 2>&1
tail -3 gpurun_out/ncu_multirhs.log
ls -la gpurun_out/prof20_multirhs.ncu-rep
