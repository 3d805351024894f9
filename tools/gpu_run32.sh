#!/bin/bash
# round-2 GPU call 32: preconditioner accumulation pass from the resident float32 encoding: tests, per-step times, fit A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fitting_gpu.py tests/test_gram_tc_gpu.py tests/test_gpu_l1.py tests/test_stream_gpu.py tests/test_config_parity_gpu.py tests/test_scoring_gpu.py -x -q -m gpu > gpurun_out/r32_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/r32_tests.log
tail -15 gpurun_out/r32_tests.log
for v in 1 0 1 0; do
  XGPR_B200_PRECOND_F64=$v timeout 600 python bench.py --no-conv --no-gram --nmll-rows 0 --no-cpu-baseline --steps 3 2>/dev/null | python -c "
import sys, json
for line in sys.stdin:
    if line.startswith('{'):
        f = json.loads(line)['fit']; print(json.dumps({'precond_f64': $v, 'fit_s': f['fit_s'], 'precond_s': f['precond_s'], 'cg_s': f['cg_s'], 'iters': f['cg_iters'], 'ratio': f['achieved_ratio'], 'werr': f['weights_vs_n1']['max_abs_err_over_max_abs_w']}))
" >> gpurun_out/r32_fit_ab.jsonl
done
cat gpurun_out/r32_fit_ab.jsonl
