mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_prefill.py -x -q > gpurun_out/pytest_prefill.log 2>&1; echo "pytest prefill rc=$?"; tail -3 gpurun_out/pytest_prefill.log
timeout 600 python tools/bench_prefill.py > gpurun_out/bench_prefill.log 2>&1; echo "bench_prefill rc=$?"; grep -v "^$" gpurun_out/bench_prefill.log | cut -c1-420
timeout 200 python tools/trace_prefill.py 0 2048 > gpurun_out/trace_prefill_s2048.log 2>&1; echo "trace rc=$?"; head -1 gpurun_out/trace_prefill_s2048.log
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -k regex:prefill --csv --log-file gpurun_out/launches_prefill_small.csv python tools/profile_prefill_small.py > gpurun_out/ncu_prefill_small.log 2>&1; echo "ncu prefill rc=$?"
timeout 300 ncu --set full --clock-control none --import-source on -k regex:prefill_attn -s 1 -c 1 -o gpurun_out/prefill_s2048_persist_r02 -f python tools/profile_prefill_small.py > gpurun_out/ncu_prefill.log 2>&1; echo "ncu prefill full rc=$?"
