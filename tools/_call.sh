mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_stream_gemm.py -x -q > gpurun_out/pytest_sgemm.log 2>&1; echo "pytest sgemm rc=$?"; tail -3 gpurun_out/pytest_sgemm.log
for s in "64 1024 8192 none" "64 8192 1280 none" "64 8192 7168 swiglu" "64 28672 8192 none"; do timeout 60 python tools/trace_stream_gemm.py $s; done > gpurun_out/sgemm_trace.log 2>&1; cat gpurun_out/sgemm_trace.log | cut -c1-1500
timeout 200 python tools/bench_stream_gemm.py --shape llama3-70b --tp 8 --batch 64 > gpurun_out/sgemm_l70tp8.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/sgemm_l70tp8.log | cut -c1-600
timeout 200 python tools/bench_stream_gemm.py --shape llama3-70b --tp 1 --batch 64 > gpurun_out/sgemm_l70tp1.log 2>&1; echo "rc=$?"; tail -5 gpurun_out/sgemm_l70tp1.log | cut -c1-500
