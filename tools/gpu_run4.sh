#!/bin/bash
# round-2 GPU call 4: tests + A/B of the experiment switches (wide E for P = 512 / 1024, conv cos/sin groups of 4, conv bulk staging)
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r4_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r4_pytest.log
for rep in 1 2; do
  XB_CASE="big" timeout 300 python tools/bench_configs.py rbf >> gpurun_out/r4_rbf_default.jsonl 2>&1
  XB_CASE="big" XGPR_B200_RBF_WIDE_E=1 timeout 300 python tools/bench_configs.py rbf >> gpurun_out/r4_rbf_wide.jsonl 2>&1
  timeout 300 python tools/bench_configs.py conv >> gpurun_out/r4_conv_default.jsonl 2>&1
  XGPR_B200_CONV_CH4=1 timeout 300 python tools/bench_configs.py conv >> gpurun_out/r4_conv_ch4.jsonl 2>&1
  XGPR_B200_CONV_NO_BULK=1 timeout 300 python tools/bench_configs.py conv >> gpurun_out/r4_conv_nobulk.jsonl 2>&1
  XGPR_B200_CONV_NO_ORDER=1 timeout 300 python tools/bench_configs.py conv >> gpurun_out/r4_conv_noorder.jsonl 2>&1
done
timeout 300 python tools/bench_configs.py misc > gpurun_out/r4_misc.jsonl 2>&1
prof() {  # name, case filter, group, kernel regex
  XB_CASE="$2" timeout 300 python tools/bench_configs.py $3 > gpurun_out/plain_$1.log 2>&1 && \
  XB_CASE="$2" timeout 600 ncu --set full --clock-control none --import-source on -k regex:$4 -s 2 -c 1 -f -o gpurun_out/prof4_$1 python tools/bench_configs.py $3 > gpurun_out/ncu_$1.log 2>&1
}
prof rbfcta "big cta" rbf rbf_cta2
prof conv "headline Conv1dRBF" conv conv_core
