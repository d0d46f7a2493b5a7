#!/bin/bash
mkdir -p gpurun_out
echo "== pytest all gpu"; timeout 1800 python -m pytest tests -m gpu -q --timeout 900 -p no:cacheprovider > gpurun_out/pytest_all.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed|AssertionError:" gpurun_out/pytest_all.log | head -30
echo "== tune l8 nopdl"; timeout 400 python tools/tune_decode.py --shape l8 --variants 6,8,9,10,11 --splits 1,2,4 --no-fa > gpurun_out/tune_l8b.log 2>&1; echo "rc=$?"; cat gpurun_out/tune_l8b.log | tail -17
echo "== tune l8 pdl"; timeout 400 python tools/tune_decode.py --shape l8 --variants 6,8,9,10,11 --splits 1,2,4 --no-fa --pdl > gpurun_out/tune_l8_pdl.log 2>&1; echo "rc=$?"; cat gpurun_out/tune_l8_pdl.log | tail -17
echo "== bench"; timeout 600 python bench.py > gpurun_out/bench.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench.log | cut -c1-1800
echo "== bench nopdl"; timeout 600 python bench.py --no-pdl --no-cpu-baseline > gpurun_out/bench_nopdl.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench_nopdl.log | cut -c1-400
echo "== prefill bench"; timeout 600 python tools/bench_prefill.py > gpurun_out/prefill_bench.log 2>&1; echo "rc=$?"; cat gpurun_out/prefill_bench.log | tail -8
echo "== elementwise bench"; timeout 600 python tools/bench_elementwise.py > gpurun_out/elementwise_bench.log 2>&1; echo "rc=$?"; grep swiglu gpurun_out/elementwise_bench.log
echo "== decode step l8"; timeout 600 python tools/bench_decode_step.py --shape llama3-8b --layers 32 --batch 32 --ctx 4096 > gpurun_out/step_l8.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/step_l8.log
