#!/bin/bash
mkdir -p gpurun_out
echo "== pytest model"; timeout 1500 python -m pytest tests/test_gpu_model.py -m gpu -q -s --timeout 900 -p no:cacheprovider > gpurun_out/pytest_model.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed|Error|^seed " gpurun_out/pytest_model.log | head -30
echo "== pytest prefill"; timeout 600 python -m pytest tests/test_gpu_prefill.py -m gpu -q --timeout 300 -p no:cacheprovider > gpurun_out/pytest_prefill.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed|AssertionError:" gpurun_out/pytest_prefill.log | head
