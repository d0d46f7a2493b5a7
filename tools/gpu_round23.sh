#!/bin/bash
# Session-13 call 11 (2 GPUs): real peer-memory tests (incl. LL push form), TP=2 check + L70 step per reduction form.
mkdir -p gpurun_out
echo "== pytest allreduce (2 GPUs)"; timeout 600 python -m pytest tests/test_gpu_allreduce_norm.py -m gpu -q --timeout 500 -p no:cacheprovider -k "peer_memory" > gpurun_out/pytest_ar23.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed" gpurun_out/pytest_ar23.log | head
echo "== bench_allreduce N=2"; timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29571 tools/bench_allreduce.py > gpurun_out/allreduce_tp2.log 2>&1; echo rc=$?; grep -E "^\{" gpurun_out/allreduce_tp2.log | cut -c1-700
for M in ll oneshot; do
echo "== check_tp N=2 $M"; B200_AR_MODE=$M timeout 200 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 tools/check_tp.py > gpurun_out/check_tp2_$M.log 2>&1; echo "rc=$?"; grep -E "^\{" gpurun_out/check_tp2_$M.log | head -3 | cut -c1-400
echo "== step l70 tp2 $M"; B200_AR_MODE=$M timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29581 tools/bench_decode_step.py --shape llama3-70b --layers 16 --batch 64 --ctx 8192 > gpurun_out/step_l70_tp2_$M.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/step_l70_tp2_$M.log | cut -c1-300
done
