#!/bin/bash
# round-2 GPU call 39: random-shape sweep of the encoded-input preconditioner steps
mkdir -p gpurun_out
timeout 1200 python tools/fuzz_precond_c32.py 80 3 > gpurun_out/r39_fuzz_precond.json 2>&1
tail -3 gpurun_out/r39_fuzz_precond.json
