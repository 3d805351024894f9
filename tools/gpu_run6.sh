#!/bin/bash
# round-2 GPU call 6: verification of the final tree -- tests, smoke, bench (N = 1), reference arm, per-config table
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q --timeout 600 > gpurun_out/r6_pytest.log 2>&1; echo "pytest rc=$?" >> gpurun_out/r6_pytest.log
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r6_smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/r6_smoke.log
timeout 600 python bench.py --steps 20 --warmup 5 > gpurun_out/r6_bench.json 2> gpurun_out/r6_bench.err; echo "bench rc=$?" >> gpurun_out/r6_bench.err
timeout 300 python bench.py --impl reference --steps 5 --warmup 2 > gpurun_out/r6_bench_ref.json 2> gpurun_out/r6_bench_ref.err
timeout 600 python tools/bench_configs.py rbf conv misc > gpurun_out/r6_bench_configs.jsonl 2> gpurun_out/r6_bench_configs.err
