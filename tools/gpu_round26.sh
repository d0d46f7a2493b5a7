#!/bin/bash
# Session-26 call 3 (1 GPU): skinny GEMM v2 (8 warps, fragments loaded up front): correctness + sweep.
mkdir -p gpurun_out
echo "== pytest skinny"; timeout 200 python -m pytest tests/test_gpu_skinny_gemm.py -m gpu -q --timeout 150 -p no:cacheprovider > gpurun_out/pytest_skinny.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed|Error|error" gpurun_out/pytest_skinny.log | head -20
echo "== sweep l8 tp1 b32"; timeout 150 python tools/bench_skinny_gemm.py --shape llama3-8b --tp 1 --batch 32 > gpurun_out/skinny_l8.log 2>&1; echo "rc=$?"; cut -c1-420 gpurun_out/skinny_l8.log | tail -5
echo "== sweep l70 tp8 b64"; timeout 150 python tools/bench_skinny_gemm.py --shape llama3-70b --tp 8 --batch 64 > gpurun_out/skinny_l70tp8.log 2>&1; echo "rc=$?"; cut -c1-420 gpurun_out/skinny_l70tp8.log | tail -5
