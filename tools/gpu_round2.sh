#!/bin/bash
mkdir -p gpurun_out
echo "== pytest all gpu (decode/elementwise/fa)"; timeout 1200 python -m pytest tests/test_gpu_elementwise.py tests/test_gpu_decode.py tests/test_gpu_vs_flash_attn.py -m gpu -q --timeout 300 -p no:cacheprovider > gpurun_out/pytest_decode.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed|AssertionError:" gpurun_out/pytest_decode.log | head -30
echo "== pytest prefill"; timeout 600 python -m pytest tests/test_gpu_prefill.py -m gpu -q --timeout 120 -p no:cacheprovider > gpurun_out/pytest_prefill.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed|AssertionError:" gpurun_out/pytest_prefill.log | head -30
echo "== golden gpu"; timeout 300 python tests/golden/make_golden_gpu.py > gpurun_out/golden_gpu.log 2>&1; echo "rc=$?"; tail -8 gpurun_out/golden_gpu.log
for sh in l8b1 l8b8 l70tp8 l70tp2 q7 long q05; do
  echo "== tune $sh"; timeout 400 python tools/tune_decode.py --shape $sh --variants 1,6,8 --splits 1,2,3,4,6,8,12,16,24,32,48,64 > gpurun_out/tune_$sh.log 2>&1; echo "rc=$?"; grep -E "best|flash-attn" gpurun_out/tune_$sh.log
done
echo "== bench"; timeout 600 python bench.py > gpurun_out/bench.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/bench.log
echo "== ncu launches"; timeout 300 python tools/profile_decode.py 6 0 1 > gpurun_out/prof_plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/launches_decode.csv python tools/profile_decode.py 6 0 1 > gpurun_out/ncu_launch.log 2>&1; echo "rc=$?"
echo "== ncu full"; timeout 300 python tools/profile_decode.py 6 0 1 > gpurun_out/prof_plain2.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:paged_decode -s 3 -c 2 -f -o gpurun_out/decode_s1 python tools/profile_decode.py 6 0 1 > gpurun_out/ncu_full.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/ncu_full.log
