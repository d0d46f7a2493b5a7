#!/bin/bash
# Session-5 call 2 (2 GPUs): fused peer-memory all-reduce + residual + RMSNorm — kernel tests, TP=2 model check,
# decode-step tokens/s fused vs NCCL.
mkdir -p gpurun_out
echo "== pytest allreduce_norm + model"; timeout 900 python -m pytest tests/test_gpu_allreduce_norm.py tests/test_gpu_model.py -m gpu -x -q --timeout 600 -p no:cacheprovider > gpurun_out/pytest_ar.log 2>&1; echo "rc=$?"; tail -15 gpurun_out/pytest_ar.log
echo "== check_tp N=2"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 tools/check_tp.py > gpurun_out/check_tp2.log 2>&1; echo "rc=$?"; tail -4 gpurun_out/check_tp2.log
for F in 1 0; do
echo "== step l70 tp2 fused=$F"; B200_FUSED_ALLREDUCE=$F timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2957$F tools/bench_decode_step.py --shape llama3-70b --layers 16 --batch 64 --ctx 8192 > gpurun_out/step_l70_tp2_fused$F.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/step_l70_tp2_fused$F.log
done
echo "== step l70 tp1"; timeout 600 python tools/bench_decode_step.py --shape llama3-70b --layers 16 --batch 64 --ctx 8192 > gpurun_out/step_l70_tp1.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/step_l70_tp1.log
