#!/bin/bash
# round-2 GPU call 42: ncu launch list of a short bench on the final tree (kernel shares of the step)
mkdir -p gpurun_out
ARGS="--steps 2 --warmup 1 --fit-rows 262144 --nmll-rows 131072 --no-cpu-baseline"
timeout 600 python bench.py $ARGS > gpurun_out/r42_plain.json 2> gpurun_out/r42_plain.err || exit 1
timeout 1500 ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/r42_launches.csv python bench.py $ARGS > gpurun_out/r42_ncu.log 2>&1
tail -2 gpurun_out/r42_ncu.log | cut -c1-300
wc -l gpurun_out/r42_launches.csv
