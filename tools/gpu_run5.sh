#!/bin/bash
# round-2 GPU call 5: tests, full bench (N = 1, overlapped fit), per-config table, ncu of conv gradient + ragged conv
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r5_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r5_pytest.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r5_bench.json 2> gpurun_out/r5_bench.err; echo "bench rc=$?" >> gpurun_out/r5_bench.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r5_bench_ref.json 2> gpurun_out/r5_bench_ref.err
timeout 600 python tools/bench_configs.py rbf conv misc > gpurun_out/r5_bench_configs.jsonl 2> gpurun_out/r5_bench_configs.err
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r5_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r5_smoke.log
prof() {  # name, case filter, group, kernel regex
  XB_CASE="$2" timeout 300 python tools/bench_configs.py $3 > gpurun_out/plain_$1.log 2>&1 && \
  XB_CASE="$2" timeout 600 ncu --set full --clock-control none --import-source on -k regex:$4 -s 2 -c 1 -f -o gpurun_out/prof5_$1 python tools/bench_configs.py $3 > gpurun_out/ncu_$1.log 2>&1
}
prof convgrad "Conv1dRBF grad" conv conv_core
prof rbf1024 "big RBF 131072x1024" rbf rbf_core
