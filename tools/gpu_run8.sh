#!/bin/bash
# round-2 GPU call 8: conv gradient with cos/sin groups of 4 (no spill), rbf_cta2 grouped epilogue -- tests + per-config table
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r8_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r8_pytest.log
for rep in 1 2; do timeout 600 python tools/bench_configs.py rbf conv >> gpurun_out/r8_bench_configs.jsonl 2>> gpurun_out/r8_bench_configs.err; done
