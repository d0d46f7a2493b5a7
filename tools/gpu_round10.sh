#!/bin/bash
# Session-5 call 3 (1 GPU): balanced (stream-K) decode kernel tests + sweeps, fused allreduce-norm single-GPU tests.
mkdir -p gpurun_out
echo "== pytest decode + allreduce_norm + model"; timeout 1500 python -m pytest tests/test_gpu_decode.py tests/test_gpu_allreduce_norm.py tests/test_gpu_model.py -m gpu -q --timeout 900 -p no:cacheprovider > gpurun_out/pytest_sk.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_sk.log | head -30
for sh in l8 l70tp8 l70tp2 l8b8 l8b16 l8b1 q7 long; do
echo "== tune $sh"; timeout 300 python tools/tune_decode.py --shape $sh --variants 6,1 --splits 0,1,2,4,8,16 --no-fa --tag _sk > gpurun_out/tune_${sh}_sk.log 2>&1; echo "rc=$?"; grep -E "splits   0|best" gpurun_out/tune_${sh}_sk.log
done
for sh in q7ragged l8ragged; do
echo "== tune $sh"; timeout 400 python tools/tune_decode.py --shape $sh --variants 6,1 --splits 0,1,2,4,8,16 --tag _sk > gpurun_out/tune_${sh}_sk.log 2>&1; echo "rc=$?"; grep -E "splits   0|best|flash" gpurun_out/tune_${sh}_sk.log
done
