#!/bin/bash
# Session-5 call (2 GPUs): LL push form of the fused all-reduce + norm; relaxed start flags.
mkdir -p gpurun_out
echo "== pytest allreduce_norm"; timeout 900 python -m pytest tests/test_gpu_allreduce_norm.py -m gpu -q --timeout 600 -p no:cacheprovider > gpurun_out/pytest_ar3.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed|Error" gpurun_out/pytest_ar3.log | head -20
echo "== bench_allreduce N=2"; timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 tools/bench_allreduce.py > gpurun_out/allreduce_tp2.log 2>&1; echo rc=$?; grep -E "^\{|Error|error" gpurun_out/allreduce_tp2.log | cut -c1-900
for M in ll oneshot; do
echo "== check_tp N=2 $M"; B200_AR_MODE=$M timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 tools/check_tp.py > gpurun_out/check_tp2_$M.log 2>&1; echo "rc=$?"; grep -E "^\{|Error|error" gpurun_out/check_tp2_$M.log | head -5
echo "== step l70 tp2 $M"; B200_AR_MODE=$M timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29581 tools/bench_decode_step.py --shape llama3-70b --layers 16 --batch 64 --ctx 8192 > gpurun_out/step_l70_tp2_$M.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/step_l70_tp2_$M.log
done
