mkdir -p gpurun_out
for c in 1 2 4 8 16; do timeout 300 python bench.py --no-decode-step --no-configs --no-cpu-baseline --e2e-chunks $c 2>&1 | grep '^{' | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('chunks', $c, 'value', j['value'], 'e2e', j['e2e']['value'], round(j['e2e']['value']/j['value'],4))"; done
B200_DECODE_SK_OV=3 timeout 300 python tools/bench_ragged.py 2>&1 | grep '^{' | cut -c1-900
timeout 900 python -m pytest tests/test_gpu_decode.py tests/test_gpu_model.py -x -q > gpurun_out/pytest_decode.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_decode.log
