#!/bin/bash
# round-2 GPU call 17: 26-column tile + float32-encoded multi-RHS matvec: tests, kernel A/B, cfg5 NMLL before / after
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fitting_gpu.py tests/test_gpu_l1.py tests/test_scoring_gpu.py tests/test_gram_tc_gpu.py tests/test_zz_no_alloc.py -x -q -m gpu > gpurun_out/r17_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/r17_tests.log
tail -5 gpurun_out/r17_tests.log
for nt in 1 0; do
  if [ $nt = 1 ]; then export XGPR_B200_MATVEC_NOTAIL=1; else unset XGPR_B200_MATVEC_NOTAIL; fi
  timeout 200 python tools/prof_multirhs_c32.py 16384 32768 26 >> gpurun_out/r17_multirhs.jsonl 2>&1
  timeout 200 python tools/prof_multirhs_c32.py 32768 16384 26 >> gpurun_out/r17_multirhs.jsonl 2>&1
done
unset XGPR_B200_MATVEC_NOTAIL
cat gpurun_out/r17_multirhs.jsonl
XGPR_B200_MATVEC_NOTAIL=1 XGPR_B200_MULTIRHS_F64=1 timeout 300 python tools/bench_nmll.py --rows 2000000 > gpurun_out/r17_nmll_before.json 2>&1
timeout 300 python tools/bench_nmll.py --rows 2000000 > gpurun_out/r17_nmll_after.json 2>&1
XGPR_B200_MULTIRHS_F64=1 timeout 300 python tools/bench_nmll.py --rows 2000000 > gpurun_out/r17_nmll_tail_only.json 2>&1
tail -2 gpurun_out/r17_nmll_before.json gpurun_out/r17_nmll_tail_only.json gpurun_out/r17_nmll_after.json
