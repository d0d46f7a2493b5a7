#!/bin/bash
# Session-26 call (1 GPU, ~13 GPU-minutes left): Triton golden vectors for the elementwise ops, full GPU
# suite on HEAD (incl. the new greedy arg-max kernel), smoke, both bench arms, sampling bench, ncu launch
# list of bench.py.  Most important first: the call may be clamped by the remaining budget.
mkdir -p gpurun_out
echo "== triton golden"; timeout 240 python tests/golden/make_golden_triton.py > gpurun_out/triton_golden_stdout.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/triton_golden_stdout.log | cut -c1-200
echo "== pytest gpu"; timeout 400 python -m pytest tests -m gpu -q --timeout 300 -p no:cacheprovider > gpurun_out/pytest_gpu24.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_gpu24.log | head -20
echo "== smoke"; timeout 120 python __graft_entry__.py smoke > gpurun_out/smoke24.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/smoke24.log | cut -c1-200
echo "== bench"; timeout 300 python bench.py > gpurun_out/bench24.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench24.log | cut -c1-1500
echo "== sampling bench"; timeout 120 python tools/bench_sampling.py > gpurun_out/sampling_bench.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/sampling_bench.log
echo "== bench ref"; timeout 300 python bench.py --impl reference > gpurun_out/bench_ref24.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench_ref24.log | cut -c1-600
echo "== ncu launch list of bench.py"; timeout 400 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_bench.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_bench24.log 2>&1; echo "rc=$?"; wc -l gpurun_out/launches_bench.csv
