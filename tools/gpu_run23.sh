#!/bin/bash
# round-2 GPU call 23: multi-RHS DMMA kernels with the conflict-free B stride and the single-barrier 2/3-stage pipeline
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fitting_gpu.py tests/test_gpu_l1.py tests/test_scoring_gpu.py tests/test_stream_gpu.py tests/test_gram_tc_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r23_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/r23_tests.log
tail -5 gpurun_out/r23_tests.log
for nt in 1 0; do
  if [ $nt = 1 ]; then export XGPR_B200_MATVEC_NOTAIL=1; else unset XGPR_B200_MATVEC_NOTAIL; fi
  timeout 200 python tools/prof_multirhs_c32.py 16384 32768 26 >> gpurun_out/r23_multirhs.jsonl 2>&1
  timeout 200 python tools/prof_multirhs_c32.py 32768 16384 26 >> gpurun_out/r23_multirhs.jsonl 2>&1
done
unset XGPR_B200_MATVEC_NOTAIL
timeout 200 python tools/prof_multirhs_c32.py 16384 32768 32 >> gpurun_out/r23_multirhs.jsonl 2>&1
cat gpurun_out/r23_multirhs.jsonl
timeout 300 python tools/bench_nmll.py --rows 2000000 > gpurun_out/r23_nmll.json 2>&1
tail -c 800 gpurun_out/r23_nmll.json
