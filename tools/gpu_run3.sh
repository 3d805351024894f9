#!/bin/bash
# round-2 GPU call 3: tests, per-config table, full bench, launch list of a short bench, ncu of the new kernels
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r3_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r3_pytest.log
timeout 600 python tools/bench_configs.py rbf conv misc > gpurun_out/r3_bench_configs.jsonl 2> gpurun_out/r3_bench_configs.err
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r3_bench.json 2> gpurun_out/r3_bench.err; echo "bench rc=$?" >> gpurun_out/r3_bench.err
timeout 300 python tools/prof_fht.py > gpurun_out/r3_fht.jsonl 2>&1
CMD="python bench.py --steps 2 --warmup 1 --fit-rows 262144 --no-cpu-baseline"
timeout 600 $CMD > gpurun_out/r3_plain_short.log 2>&1 && \
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 900 --csv --log-file gpurun_out/r3_launches.csv $CMD > gpurun_out/r3_ncu_launches.log 2>&1
prof() {  # name, case filter, group, kernel regex
  XB_CASE="$2" timeout 300 python tools/bench_configs.py $3 > gpurun_out/plain_$1.log 2>&1 && \
  XB_CASE="$2" timeout 600 ncu --set full --clock-control none --import-source on -k regex:$4 -s 2 -c 1 -f -o gpurun_out/prof3_$1 python tools/bench_configs.py $3 > gpurun_out/ncu_$1.log 2>&1
}
prof rbf512 "big cfg1-shape" rbf rbf_core
prof rbfcta "big cta" rbf rbf_cta2
prof ard "MiniARD grad 65536" misc ard_mma
