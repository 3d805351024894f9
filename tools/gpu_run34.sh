#!/bin/bash
# round-2 GPU call 34: encoded SRHT sample with the prefetch kept in the stored type: tests, per-step times, fit A/B
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_fitting_gpu.py tests/test_gpu_l1.py tests/test_stream_gpu.py tests/test_scoring_gpu.py -x -q -m gpu > gpurun_out/r34_tests.log 2>&1
echo "tests exit $?" >> gpurun_out/r34_tests.log
tail -4 gpurun_out/r34_tests.log
timeout 300 python tools/prof_precond_chunk.py 32768 16384 > gpurun_out/r34_precond_chunk.jsonl 2>&1
cat gpurun_out/r34_precond_chunk.jsonl
for v in 1 0 1 0; do
  XGPR_B200_PRECOND_F64=$v timeout 600 python bench.py --no-conv --no-gram --nmll-rows 0 --no-cpu-baseline --steps 3 2>/dev/null | python -c "
import sys, json
for line in sys.stdin:
    if line.startswith('{'):
        f = json.loads(line)['fit']; print(json.dumps({'precond_f64': $v, 'fit_s': f['fit_s'], 'precond_s': f['precond_s'], 'cg_s': f['cg_s'], 'iters': f['cg_iters'], 'ratio': f['achieved_ratio'], 'werr': f['weights_vs_n1']['max_abs_err_over_max_abs_w']}))
" >> gpurun_out/r34_fit_ab.jsonl
done
cat gpurun_out/r34_fit_ab.jsonl
