#!/bin/bash
# round-2 GPU call 11: A/B of the mixed XU / ALU widening in the resident matvec (+ parity tests under the switch)
mkdir -p gpurun_out
for rep in 1 2 3; do
  timeout 200 python tools/prof_cres.py 262144 16384 >> gpurun_out/r11_cres_default.jsonl 2>&1
  XGPR_B200_CRES_MIX=1 timeout 200 python tools/prof_cres.py 262144 16384 >> gpurun_out/r11_cres_mix.jsonl 2>&1
done
timeout 200 python tools/prof_cres.py 262144 8192 >> gpurun_out/r11_cres_default.jsonl 2>&1
XGPR_B200_CRES_MIX=1 timeout 200 python tools/prof_cres.py 262144 8192 >> gpurun_out/r11_cres_mix.jsonl 2>&1
timeout 200 python tools/prof_cres.py 262144 4096 >> gpurun_out/r11_cres_default.jsonl 2>&1
XGPR_B200_CRES_MIX=1 timeout 200 python tools/prof_cres.py 262144 4096 >> gpurun_out/r11_cres_mix.jsonl 2>&1
XGPR_B200_CRES_MIX=1 timeout 900 python -m pytest tests/test_stream_gpu.py tests/test_fitting_gpu.py tests/test_config_parity_gpu.py -m gpu -q --timeout 600 > gpurun_out/r11_pytest_mix.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r11_pytest_mix.log
XGPR_B200_CRES_MIX=1 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r11_smoke_mix.log 2>&1
XGPR_B200_CRES_MIX=1 timeout 300 python tools/prof_cres.py 262144 16384 > gpurun_out/plain_cres.log 2>&1 && \
XGPR_B200_CRES_MIX=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:cres_zzv -s 3 -c 1 -f -o gpurun_out/prof11_cres python tools/prof_cres.py 262144 16384 > gpurun_out/ncu_cres.log 2>&1
