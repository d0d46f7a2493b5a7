#!/bin/bash
# Session-5 call 1 (1 GPU): full GPU test suite, bench, per-kernel layer breakdown at TP-local shapes,
# then ONE ncu pass (--set full) over the elementwise + prefill kernels.
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 -p no:cacheprovider > gpurun_out/pytest_gpu.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/pytest_gpu.log
echo "== bench"; timeout 600 python bench.py > gpurun_out/bench.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench.log | cut -c1-600
echo "== bench reference arm"; timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench_ref.log | cut -c1-600
for tp in 8 1; do
echo "== layer l70 tp$tp"; timeout 300 python tools/profile_layer.py --shape llama3-70b --tp $tp --layers 4 --batch 64 --ctx 8192 > gpurun_out/layer_l70_tp$tp.log 2>&1; echo "rc=$?"; tail -18 gpurun_out/layer_l70_tp$tp.log
done
echo "== layer l8 tp1"; timeout 300 python tools/profile_layer.py --shape llama3-8b --tp 1 --layers 4 --batch 32 --ctx 4096 > gpurun_out/layer_l8_tp1.log 2>&1; echo "rc=$?"; tail -18 gpurun_out/layer_l8_tp1.log
echo "== ncu kernels"
timeout 300 python tools/profile_kernels.py > gpurun_out/profile_kernels_plain.log 2>&1 && \
timeout 900 ncu --set full --clock-control none --import-source on -k regex:"swiglu|rmsnorm|store_kvcache|prefill" -c 12 -f -o gpurun_out/kernels_r01b python tools/profile_kernels.py > gpurun_out/ncu_kernels.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/ncu_kernels.log
