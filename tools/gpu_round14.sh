#!/bin/bash
# Session-13 call 2 (1 GPU): prefill softmax arithmetic variants (packed fp32x2, polynomial exp2).
mkdir -p gpurun_out
echo "== pytest prefill"; timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q --timeout 600 -p no:cacheprovider > gpurun_out/pytest_pf14.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_pf14.log
echo "== bench prefill"; timeout 600 python tools/bench_prefill.py > gpurun_out/bench_prefill14.log 2>&1; echo "rc=$?"; cat gpurun_out/bench_prefill14.log | tail -8
