#!/bin/bash
# round-2 GPU call 31: random-shape sweep of the consumers with guard bands
mkdir -p gpurun_out
timeout 900 python tools/fuzz_consumers.py 300 1 > gpurun_out/r31_fuzz.json 2>&1
tail -3 gpurun_out/r31_fuzz.json
XGPR_B200_MATVEC_NOTAIL=1 timeout 600 python tools/fuzz_consumers.py 100 2 >> gpurun_out/r31_fuzz.json 2>&1
tail -1 gpurun_out/r31_fuzz.json
