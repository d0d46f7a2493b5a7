#!/bin/bash
# First GPU pass: smoke, parity tests, tuning sweep, bench.  Everything logs into gpurun_out/.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm,clocks.sm,power.limit --format=csv > gpurun_out/gpu.txt 2>&1
nproc >> gpurun_out/gpu.txt
echo "== smoke" ; timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -5 gpurun_out/smoke.log
echo "== pytest elementwise+decode"; timeout 900 python -m pytest tests/test_gpu_elementwise.py tests/test_gpu_decode.py -m gpu -q --timeout 300 -p no:cacheprovider > gpurun_out/pytest_decode.log 2>&1; echo "rc=$?"; tail -30 gpurun_out/pytest_decode.log
echo "== pytest prefill"; timeout 600 python -m pytest tests/test_gpu_prefill.py -m gpu -q --timeout 120 -x -p no:cacheprovider > gpurun_out/pytest_prefill.log 2>&1; echo "rc=$?"; tail -30 gpurun_out/pytest_prefill.log
echo "== golden gpu"; timeout 300 python tests/golden/make_golden_gpu.py > gpurun_out/golden_gpu.log 2>&1; echo "rc=$?"; tail -12 gpurun_out/golden_gpu.log
echo "== pytest vs flash-attn"; timeout 600 python -m pytest tests/test_gpu_vs_flash_attn.py -m gpu -q --timeout 120 -p no:cacheprovider > gpurun_out/pytest_fa.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/pytest_fa.log
echo "== tune l8"; timeout 600 python tools/tune_decode.py --shape l8 > gpurun_out/tune_l8.log 2>&1; echo "rc=$?"; tail -12 gpurun_out/tune_l8.log
echo "== bench"; timeout 600 python bench.py > gpurun_out/bench.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/bench.log
