mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 tools/bench_allreduce.py 64,8 > gpurun_out/allreduce_tp2.log 2>&1; echo "ar rc=$?"; grep '^{' gpurun_out/allreduce_tp2.log | cut -c1-3000
run() { name=$1; shift; env "$@" timeout 300 python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29575 tools/bench_decode_step.py --shape llama3-70b --layers 16 --batch 64 --ctx 8192 > gpurun_out/step_l70_tp2_$name.log 2>&1; echo "step $name rc=$?"; grep '^{' gpurun_out/step_l70_tp2_$name.log | python -c "import sys,json; [print({k:d[k] for k in ('ms_per_step_run','tokens_per_s_run','allreduce','gemm','ar_prefetch_MB')}) for d in map(json.loads, sys.stdin)]"; }
run nvls1 B200_AR_PREFETCH_MB=0 B200_AR_MODE=nvls1
run nvls1_pf B200_AR_MODE=nvls1
run oneshot B200_AR_PREFETCH_MB=0 B200_AR_MODE=oneshot
run nvls1_dpdl B200_AR_PREFETCH_MB=0 B200_AR_MODE=nvls1 B200_DECODE_PDL=1 B200_STREAM_GEMM_PDL=1
for s in 4 5 6; do timeout 120 python tools/trace_decode.py --shape l70tp8 --splits $s >> gpurun_out/trace_decode.log 2>&1; done; timeout 120 python tools/trace_decode.py --shape l70tp8 --splits 4 --variant 1 >> gpurun_out/trace_decode.log 2>&1; timeout 120 python tools/trace_decode.py --shape l8 >> gpurun_out/trace_decode.log 2>&1; grep '^{' gpurun_out/trace_decode.log | cut -c1-1800
