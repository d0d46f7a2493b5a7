#!/bin/bash
mkdir -p gpurun_out
echo "== pytest prefill+fa+pdl"; timeout 1200 python -m pytest tests/test_gpu_prefill.py tests/test_gpu_vs_flash_attn.py tests/test_gpu_decode.py -m gpu -q --timeout 600 -p no:cacheprovider -k "prefill or flash or dependent" > gpurun_out/pytest_p.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed|AssertionError:|Error" gpurun_out/pytest_p.log | head -30
echo "== prefill bench"; timeout 600 python tools/bench_prefill.py > gpurun_out/prefill_bench.log 2>&1; echo "rc=$?"; cat gpurun_out/prefill_bench.log | tail -8
echo "== tune l8 pdl2"; timeout 400 python tools/tune_decode.py --shape l8 --variants 6,8 --splits 1,4 --no-fa --pdl 2 > gpurun_out/tune_l8_pdl2.log 2>&1; echo "rc=$?"; cat gpurun_out/tune_l8_pdl2.log | tail -6
echo "== bench"; timeout 600 python bench.py > gpurun_out/bench.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench.log | cut -c1-1800
