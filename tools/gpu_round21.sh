#!/bin/bash
# Session-13 call 9 (1 GPU): full suite with the pair kernel as default prefill path, prefill bench, ncu capture.
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -m gpu -q --timeout 900 -p no:cacheprovider > gpurun_out/pytest_gpu21.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu21.log | head
echo "== smoke"; timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke21.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/smoke21.log
echo "== bench prefill"; timeout 600 python tools/bench_prefill.py > gpurun_out/bench_prefill21.log 2>&1; echo "rc=$?"; tail -6 gpurun_out/bench_prefill21.log | cut -c1-330
timeout 300 python tools/profile_kernels.py prefill > gpurun_out/prof_prefill_plain.log 2>&1; echo "plain rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:prefill_attn -s 1 -c 1 -f -o gpurun_out/prefill_r01d python tools/profile_kernels.py prefill > gpurun_out/ncu_prefill.log 2>&1; echo "ncu rc=$?"
