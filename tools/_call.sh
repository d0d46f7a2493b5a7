mkdir -p gpurun_out
W=${WORLD:-8}
timeout 300 python -m torch.distributed.run --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29571 tools/bench_allreduce.py 64 > gpurun_out/allreduce_tp$W.log 2>&1; echo "ar rc=$?"; grep '^{' gpurun_out/allreduce_tp$W.log | cut -c1-3500
run() { name=$1; shift; env "$@" timeout 240 python -m torch.distributed.run --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29575 tools/bench_decode_step.py --shape llama3-70b --layers 16 --batch 64 --ctx 8192 $EXTRA > gpurun_out/step_l70_tp${W}_$name.log 2>&1; echo "step $name rc=$?"; grep '^{' gpurun_out/step_l70_tp${W}_$name.log | python -c "import sys,json; [print({k:d[k] for k in ('ms_per_step_run','tokens_per_s_run','allreduce','gemm','ar_prefetch_MB')}) for d in map(json.loads, sys.stdin)]"; }
run nvls1 B200_AR_PREFETCH_MB=0
run nvls1_pf48 B200_AR_PREFETCH_MB=48
run nvls B200_AR_PREFETCH_MB=0 B200_AR_MODE=nvls
run ll B200_AR_PREFETCH_MB=0 B200_AR_MODE=ll
run oneshot B200_AR_PREFETCH_MB=0 B200_AR_MODE=oneshot
run nvls1_gpdl B200_AR_PREFETCH_MB=0 B200_STREAM_GEMM_PDL=1
run nvls1_gpdl_dpdl B200_AR_PREFETCH_MB=0 B200_STREAM_GEMM_PDL=1 B200_DECODE_PDL=1 B200_AR_PDL=1
run nvls1_cublas B200_AR_PREFETCH_MB=0 B200_STREAM_GEMM=0
run nvls1_allstream B200_AR_PREFETCH_MB=0 B200_STREAM_GEMM=1 B200_STREAM_GEMM_PDL=1
EXTRA=--profile run nvls1_profile B200_AR_PREFETCH_MB=0
ls gpurun_out | head -40
