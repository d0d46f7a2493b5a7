#!/bin/bash
# Session-5 call 4 (2 GPUs): row-sharded fused all-reduce + norm; peer-memory tests; TP=2 checks both modes.
mkdir -p gpurun_out
echo "== pytest allreduce_norm + decode(balanced)"; timeout 900 python -m pytest tests/test_gpu_allreduce_norm.py tests/test_gpu_decode.py -m gpu -q --timeout 600 -p no:cacheprovider -k "allreduce or balanced or sharded or peer or world1 or emulated or flag" > gpurun_out/pytest_ar2.log 2>&1; echo "rc=$?"; grep -E "^(FAILED|ERROR)|passed|failed|Error" gpurun_out/pytest_ar2.log | head -20
for M in 1 0; do
echo "== check_tp N=2 sharded=$M"; B200_AR_SHARDED=$M timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2956$M tools/check_tp.py > gpurun_out/check_tp2_sh$M.log 2>&1; echo "rc=$?"; grep -E "^\{|Error|error" gpurun_out/check_tp2_sh$M.log | head -5
echo "== step l70 tp2 sharded=$M"; B200_AR_SHARDED=$M timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2958$M tools/bench_decode_step.py --shape llama3-70b --layers 16 --batch 64 --ctx 8192 > gpurun_out/step_l70_tp2_sh$M.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/step_l70_tp2_sh$M.log
done
