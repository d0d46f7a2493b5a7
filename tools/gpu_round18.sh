#!/bin/bash
# Session-13 call 6 (1 GPU): two-Q-tiles-per-CTA prefill kernel (variants 6-8): tests, bench, trace.
mkdir -p gpurun_out
echo "== pytest prefill"; timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -q --timeout 600 -p no:cacheprovider > gpurun_out/pytest_pf18.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_pf18.log | head -30
echo "== bench prefill"; timeout 600 python tools/bench_prefill.py > gpurun_out/bench_prefill18.log 2>&1; echo "rc=$?"; cat gpurun_out/bench_prefill18.log | tail -8
for v in 6; do timeout 300 python tools/trace_prefill.py $v 8192 > gpurun_out/trace_v$v.log 2>&1; echo "v$v rc=$?"; head -1 gpurun_out/trace_v$v.log; done
