#!/bin/bash
# round-2 GPU call 36: cfg5 NMLL with a larger share of HBM for the kept encoding (0.75 / 0.85 / 0.9), same box
mkdir -p gpurun_out
for f in 0.75 0.85 0.9 0.75 0.85; do
  XGPR_B200_ENCODED_FRACTION=$f timeout 300 python tools/bench_nmll.py --rows 2000000 2>&1 | grep '^{' | python -c "
import sys, json
d = json.loads(sys.stdin.read()); print(json.dumps({'fraction': $f, 'seconds': d['seconds'], 'peak_mem_gb': d['peak_mem_gb'], **d['phases_rank0']}))" >> gpurun_out/r36_nmll_fraction.jsonl
done
cat gpurun_out/r36_nmll_fraction.jsonl
