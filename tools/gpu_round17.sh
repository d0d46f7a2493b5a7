#!/bin/bash
# Session-13 call 5 (1 GPU): warp-uniform MMA issue (elect.sync) in the prefill kernel: tests, bench, trace.
mkdir -p gpurun_out
echo "== pytest prefill"; timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -x -q --timeout 600 -p no:cacheprovider > gpurun_out/pytest_pf17.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/pytest_pf17.log
echo "== bench prefill"; timeout 600 python tools/bench_prefill.py > gpurun_out/bench_prefill17.log 2>&1; echo "rc=$?"; cat gpurun_out/bench_prefill17.log | tail -8
for v in 2 3 5; do timeout 300 python tools/trace_prefill.py $v 8192 > gpurun_out/trace_v$v.log 2>&1; echo "v$v rc=$?"; head -1 gpurun_out/trace_v$v.log; done
