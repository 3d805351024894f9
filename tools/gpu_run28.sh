#!/bin/bash
# round-2 GPU call 28: guard-band test of the consumers' outputs + the whole GPU suite on the last tree
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/r28_pytest_gpu.log 2>&1
echo "suite exit $?" >> gpurun_out/r28_pytest_gpu.log
tail -6 gpurun_out/r28_pytest_gpu.log
