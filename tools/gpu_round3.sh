#!/bin/bash
mkdir -p gpurun_out
echo "== pytest model"; timeout 1500 python -m pytest tests/test_gpu_model.py -m gpu -q -s --timeout 900 -p no:cacheprovider > gpurun_out/pytest_model.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed|Error|seed " gpurun_out/pytest_model.log | head -30
echo "== prefill bench"; timeout 600 python tools/bench_prefill.py > gpurun_out/prefill_bench.log 2>&1; echo "rc=$?"; cat gpurun_out/prefill_bench.log | tail -8
echo "== elementwise bench"; timeout 600 python tools/bench_elementwise.py > gpurun_out/elementwise_bench.log 2>&1; echo "rc=$?"; tail -50 gpurun_out/elementwise_bench.log
echo "== decode step l8"; timeout 600 python tools/bench_decode_step.py --shape llama3-8b --layers 32 --batch 32 --ctx 4096 > gpurun_out/step_l8.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/step_l8.log
echo "== decode step l70 tp1"; timeout 900 python tools/bench_decode_step.py --shape llama3-70b --layers 16 --batch 64 --ctx 8192 > gpurun_out/step_l70_tp1.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/step_l70_tp1.log
