#!/bin/bash
# Session-26 call 4 (1 GPU): ncu --set full of the skinny GEMM (what bounds it), then the full GPU suite, smoke and bench on HEAD.
mkdir -p gpurun_out
echo "== run profile_skinny plain"; timeout 120 python tools/profile_skinny.py > gpurun_out/profile_skinny.log 2>&1; echo "rc=$?"
echo "== ncu skinny"; timeout 300 ncu --set full --clock-control none --import-source on -k regex:skinny_gemm -c 6 -o gpurun_out/skinny_full -f python tools/profile_skinny.py > gpurun_out/ncu_skinny.log 2>&1; echo "rc=$?"; ls -la gpurun_out/skinny_full.ncu-rep
echo "== pytest gpu"; timeout 400 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider > gpurun_out/pytest_gpu27.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu27.log | head -20
echo "== smoke"; timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke27.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/smoke27.log | cut -c1-200
echo "== bench"; timeout 300 python bench.py > gpurun_out/bench27.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench27.log | cut -c1-400
