#!/bin/bash
# round-2 GPU call 25: 16-byte-load single-RHS Z^T y / Z^T (Z v): tests + A/B + fit
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fitting_gpu.py tests/test_gpu_l1.py tests/test_scoring_gpu.py tests/test_stream_gpu.py tests/test_zz_no_alloc.py tests/test_config_parity_gpu.py -x -q -m gpu > gpurun_out/r25_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/r25_tests.log
tail -4 gpurun_out/r25_tests.log
for v in 1 0 1 0; do
  if [ $v = 1 ]; then export XGPR_B200_ZTY_V1=1; else unset XGPR_B200_ZTY_V1; fi
  timeout 200 python tools/prof_zty.py 32768 16384 >> gpurun_out/r25_zty.jsonl 2>&1
  timeout 200 python tools/prof_zty.py 16384 32768 >> gpurun_out/r25_zty.jsonl 2>&1
done
unset XGPR_B200_ZTY_V1
timeout 200 python tools/prof_zty.py 2000 16384 >> gpurun_out/r25_zty.jsonl 2>&1
cat gpurun_out/r25_zty.jsonl
for v in 1 0; do
  if [ $v = 1 ]; then export XGPR_B200_ZTY_V1=1; else unset XGPR_B200_ZTY_V1; fi
  timeout 600 python bench.py --no-conv --no-gram --nmll-rows 0 --no-cpu-baseline --steps 3 2>/dev/null | python -c "
import sys, json
for line in sys.stdin:
    if line.startswith('{'):
        f = json.loads(line)['fit']; print(json.dumps({'zty_v1': $v, 'fit_s': f['fit_s'], 'precond_s': f['precond_s'], 'cg_s': f['cg_s']}))
" >> gpurun_out/r25_fit_ab.jsonl
done
cat gpurun_out/r25_fit_ab.jsonl
