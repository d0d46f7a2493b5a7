#!/bin/bash
# Session-13 call 7 (1 GPU): pair kernel, single-pass softmax with the score row in registers.
mkdir -p gpurun_out
echo "== pytest prefill"; timeout 900 python -m pytest tests/test_gpu_prefill.py -m gpu -q --timeout 600 -p no:cacheprovider -k "two_tiles" > gpurun_out/pytest_pf19.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_pf19.log | head -30
echo "== bench prefill"; timeout 600 python tools/bench_prefill.py > gpurun_out/bench_prefill19.log 2>&1; echo "rc=$?"; cat gpurun_out/bench_prefill19.log | tail -8
for v in 6; do timeout 300 python tools/trace_prefill.py $v 8192 > gpurun_out/trace_v$v.log 2>&1; echo "v$v rc=$?"; head -1 gpurun_out/trace_v$v.log; done
