#!/bin/bash
# round-2 GPU call 33: per-step times of the preconditioner chunk, fp64 vs encoded input
mkdir -p gpurun_out
timeout 300 python tools/prof_precond_chunk.py 32768 16384 > gpurun_out/r33_precond_chunk.jsonl 2>&1
cat gpurun_out/r33_precond_chunk.jsonl
