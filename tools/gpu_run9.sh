#!/bin/bash
# round-2 GPU call 9: MiniARD gradient with 16-byte epilogue accesses -- tests + timing + ncu
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r9_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r9_pytest.log
timeout 300 python tools/bench_configs.py misc > gpurun_out/r9_misc.jsonl 2> gpurun_out/r9_misc.err
XB_CASE="MiniARD grad 65536" timeout 300 python tools/bench_configs.py misc > gpurun_out/plain_ard9.log 2>&1 && \
XB_CASE="MiniARD grad 65536" timeout 600 ncu --set full --clock-control none --import-source on -k regex:ard_mma -s 2 -c 1 -f -o gpurun_out/prof9_ard python tools/bench_configs.py misc > gpurun_out/ncu_ard9.log 2>&1
