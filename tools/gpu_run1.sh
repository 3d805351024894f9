#!/bin/bash
# round-2 GPU call 1: tests, bench (+ N=1 fit weights), f32 parity study, per-config table, ncu captures
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/r1_smi.txt 2>&1
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r1_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r1_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 --save-fit-weights > gpurun_out/r1_bench.json 2> gpurun_out/r1_bench.err; echo "bench rc=$?" >> gpurun_out/r1_bench.err
cp tests/golden/cfg2_fit_w_n1.npy gpurun_out/ 2>/dev/null
timeout 300 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r1_bench_ref.json 2> gpurun_out/r1_bench_ref.err
timeout 600 python tools/f32_parity_study.py > gpurun_out/r1_f32_parity.json 2> gpurun_out/r1_f32.err; echo "f32 rc=$?" >> gpurun_out/r1_f32.err
timeout 600 python tools/bench_configs.py rbf conv misc > gpurun_out/r1_bench_configs.jsonl 2> gpurun_out/r1_bench_configs.err
prof() {  # name, case filter, group, kernel regex
  XB_CASE="$2" timeout 300 python tools/bench_configs.py $3 > gpurun_out/plain_$1.log 2>&1 && \
  XB_CASE="$2" timeout 600 ncu --set full --clock-control none --import-source on -k regex:$4 -s 2 -c 1 -f -o gpurun_out/prof_$1 python tools/bench_configs.py $3 > gpurun_out/ncu_$1.log 2>&1
}
prof rbf512 "big cfg1-shape" rbf rbf_core
prof rbf1024 "big RBF 131072x1024" rbf rbf_core
prof rbfcta "big cta" rbf rbf_cta
prof ard "MiniARD grad 65536" misc ard_grad
prof convgrad "Conv1dRBF grad" conv conv_core
ls -la gpurun_out > gpurun_out/r1_ls.txt
