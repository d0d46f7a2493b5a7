mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_allreduce_norm.py -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 tools/bench_allreduce.py 64 > gpurun_out/allreduce_tp2.log 2>&1; echo "ar rc=$?"; grep '^{' gpurun_out/allreduce_tp2.log | cut -c1-3000
