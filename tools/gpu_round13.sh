#!/bin/bash
# Session-13 call (1 GPU): full -m gpu suite, smoke, both bench arms on HEAD.
mkdir -p gpurun_out
echo "== pytest -m gpu"; timeout 1500 python -m pytest tests -m gpu -x -q --timeout 900 -p no:cacheprovider > gpurun_out/pytest_gpu13.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/pytest_gpu13.log
echo "== smoke"; timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke13.log 2>&1; echo "rc=$?"; tail -2 gpurun_out/smoke13.log
echo "== bench"; timeout 600 python bench.py > gpurun_out/bench13.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench13.log | cut -c1-1500
echo "== bench ref"; timeout 600 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_ref13.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench_ref13.log | cut -c1-600
