#!/bin/bash
# round-2 GPU call 21: tc_gram_pair_kernel (cta_group::2) -- parity tests under a timeout, then the Gram A/B
mkdir -p gpurun_out
XGPR_B200_TC_PAIR=1 timeout 300 python -m pytest tests/test_gram_tc_gpu.py -x -q -m gpu > gpurun_out/r21_pair_tests.log 2>&1
echo "pair tests exit $?" >> gpurun_out/r21_pair_tests.log
tail -15 gpurun_out/r21_pair_tests.log
if grep -q "pair tests exit 0" gpurun_out/r21_pair_tests.log; then
  for v in 0 1 0 1; do
    XGPR_B200_TC_PAIR=$v timeout 300 python bench.py --no-conv --fit-rows 0 --nmll-rows 0 --no-cpu-baseline --steps 3 2>/dev/null | python -c "
import sys, json
for line in sys.stdin:
    if line.startswith('{'):
        d = json.loads(line); g = d['gram']; print(json.dumps({'pair': $v, 'tflops': g['value'], 'ms_per_call': g['ms_per_call'], 'gemm_ms': g.get('gemm_ms'), 'roofline': g.get('roofline')}))
" >> gpurun_out/r21_gram_ab.jsonl
  done
  cat gpurun_out/r21_gram_ab.jsonl
  XGPR_B200_TC_PAIR=1 timeout 600 python -m pytest tests/test_fitting_gpu.py tests/test_scoring_gpu.py tests/test_config_parity_gpu.py -x -q -m gpu > gpurun_out/r21_pair_fit_tests.log 2>&1
  echo "pair fit tests exit $?" >> gpurun_out/r21_pair_fit_tests.log
  tail -4 gpurun_out/r21_pair_fit_tests.log
fi
