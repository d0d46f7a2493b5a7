#!/bin/bash
# Session-26 final call (1 GPU): full GPU suite, smoke, bench on HEAD with the skinny GEMM policy on by default;
# decode step of the Llama-3-8B shape with the policy on / off.
mkdir -p gpurun_out
echo "== pytest gpu"; timeout 400 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider > gpurun_out/pytest_gpu28.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu28.log | head -20
for F in auto 0; do
echo "== step l8 skinny=$F"; B200_SKINNY_GEMM=$F timeout 150 python tools/bench_decode_step.py --shape llama3-8b --layers 32 --batch 32 --ctx 4096 > gpurun_out/step_l8_skinny_$F.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/step_l8_skinny_$F.log | cut -c1-300
done
echo "== smoke"; timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke28.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/smoke28.log | cut -c1-200
echo "== bench"; timeout 300 python bench.py > gpurun_out/bench28.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench28.log | cut -c1-300
