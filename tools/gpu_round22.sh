#!/bin/bash
# Session-13 call 10 (1 GPU): LPT grid order for the prefill kernels.
mkdir -p gpurun_out
echo "== pytest prefill+model"; timeout 1200 python -m pytest tests/test_gpu_prefill.py tests/test_gpu_model.py -m gpu -q --timeout 900 -p no:cacheprovider > gpurun_out/pytest_pf22.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_pf22.log | head
echo "== bench prefill"; timeout 600 python tools/bench_prefill.py > gpurun_out/bench_prefill22.log 2>&1; echo "rc=$?"; grep -oE "'case': '[a-z0-9_]+'|'tflops': [0-9.]+|'graph_tflops': [0-9.]+|'sm_poly25_tflops': [0-9.]+|'pair_poly25_tflops': [0-9.]+" gpurun_out/bench_prefill22.log | paste - - - - -
