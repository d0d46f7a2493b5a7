#!/bin/bash
# small-batch check of the skinny GEMM policy (M = 1, 8, 16: one 16-row tile) against cuBLAS, quick sweep
mkdir -p gpurun_out
for B in 1 8 16; do
timeout 60 python tools/bench_skinny_gemm.py --shape llama3-8b --tp 1 --batch $B --quick > gpurun_out/skinny_l8_b$B.log 2>&1; echo "b=$B rc=$?"; grep '"down"' gpurun_out/skinny_l8_b$B.log | cut -c1-400
done
