mkdir -p gpurun_out
timeout 300 python tools/_emul_long.py 2>&1 | tail -4
timeout 600 python bench.py --no-decode-step --no-cpu-baseline > gpurun_out/bench_cfg.log 2>&1; echo "bench rc=$?"; tail -1 gpurun_out/bench_cfg.log | python -c "import sys,json; j=json.loads(sys.stdin.read()); print(json.dumps(j['configs'],indent=0)[:3000])"
