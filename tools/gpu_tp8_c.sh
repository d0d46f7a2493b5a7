#!/bin/bash
# 8 GPUs: reduction microbench (one-shot / sharded / LL / NCCL) + L70 decode step per reduction form.
mkdir -p gpurun_out
echo "== bench_allreduce N=8"; timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29571 tools/bench_allreduce.py 64,8 > gpurun_out/allreduce_tp8.log 2>&1; echo rc=$?; grep -E "^\{|Error|error" gpurun_out/allreduce_tp8.log | cut -c1-900
ARGS="--shape llama3-70b --layers 16 --batch 64 --ctx 8192"
for M in ll sharded oneshot; do
echo "== step l70 tp8 $M"; B200_AR_MODE=$M timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611 tools/bench_decode_step.py $ARGS > gpurun_out/step_l70_tp8_$M.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/step_l70_tp8_$M.log | cut -c1-300
done
echo "== step l70 tp4 oneshot"; B200_AR_MODE=oneshot timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29612 tools/bench_decode_step.py $ARGS > gpurun_out/step_l70_tp4_oneshot.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/step_l70_tp4_oneshot.log | cut -c1-300
