mkdir -p gpurun_out
python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 tools/bench_allreduce.py 64,8,1 > gpurun_out/allreduce_tp2.log 2>&1; echo "ar rc=$?"; grep '^{' gpurun_out/allreduce_tp2.log | cut -c1-2500
timeout 300 python -m pytest tests/test_gpu_allreduce_norm.py -x -q > gpurun_out/pytest_ar.log 2>&1; echo "pytest ar rc=$?"; tail -4 gpurun_out/pytest_ar.log
for pf in 0 48; do
B200_AR_PREFETCH_MB=$pf timeout 300 python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29575 tools/bench_decode_step.py --shape llama3-70b --layers 16 --batch 64 --ctx 8192 > gpurun_out/step_l70_tp2_pf$pf.log 2>&1; echo "step pf=$pf rc=$?"; grep '^{' gpurun_out/step_l70_tp2_pf$pf.log | cut -c1-600
done
