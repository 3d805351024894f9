#!/bin/bash
# round-2 GPU call 2: tests (incl. compat drop-in, DMMA MiniARD gradient, rbf_cta2, linear tiles), per-config table, f32 study at 3 levels
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r2_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r2_pytest.log
timeout 600 python tools/bench_configs.py rbf conv misc > gpurun_out/r2_bench_configs.jsonl 2> gpurun_out/r2_bench_configs.err
timeout 600 python tools/f32_parity_study.py > gpurun_out/r2_f32_parity.json 2> gpurun_out/r2_f32.err; echo "f32 rc=$?" >> gpurun_out/r2_f32.err
timeout 600 python bench.py --steps 10 --warmup 3 --no-gram --fit-rows 0 > gpurun_out/r2_bench.json 2> gpurun_out/r2_bench.err; echo "bench rc=$?" >> gpurun_out/r2_bench.err
prof() {  # name, case filter, group, kernel regex
  XB_CASE="$2" timeout 300 python tools/bench_configs.py $3 > gpurun_out/plain_$1.log 2>&1 && \
  XB_CASE="$2" timeout 600 ncu --set full --clock-control none --import-source on -k regex:$4 -s 2 -c 1 -f -o gpurun_out/prof2_$1 python tools/bench_configs.py $3 > gpurun_out/ncu_$1.log 2>&1
}
prof rbf512 "big cfg1-shape" rbf rbf_core
prof rbfcta "big cta" rbf rbf_cta2
prof ard "MiniARD grad 65536" misc ard_mma
