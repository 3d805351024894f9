#!/bin/bash
# round-2 GPU call 7: conv gradient at 2 CTAs/SM, refreshed ncu evidence for the final tree (launch list of a short
# bench, --set full of the dominant kernel), tests
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r7_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r7_pytest.log
timeout 600 python tools/bench_configs.py conv misc > gpurun_out/r7_bench_configs.jsonl 2> gpurun_out/r7_bench_configs.err
CMD="python bench.py --steps 2 --warmup 1 --fit-rows 262144 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/r7_plain_short.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 1200 --csv --log-file gpurun_out/r7_launches.csv $CMD > gpurun_out/r7_ncu_launches.log 2>&1
prof() {  # name, case filter, group, kernel regex
  XB_CASE="$2" timeout 300 python tools/bench_configs.py $3 > gpurun_out/plain_$1.log 2>&1 && \
  XB_CASE="$2" timeout 600 ncu --set full --clock-control none --import-source on -k regex:$4 -s 2 -c 1 -f -o gpurun_out/prof7_$1 python tools/bench_configs.py $3 > gpurun_out/ncu_$1.log 2>&1
}
prof rbf64 "cfg2 RBF" rbf rbf_core
prof convgrad "Conv1dRBF grad" conv conv_core
