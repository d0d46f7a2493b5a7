#!/bin/bash
mkdir -p gpurun_out
for N in 8 4; do
echo "== step l70 tp$N"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$N tools/bench_decode_step.py --shape llama3-70b --layers 16 --batch 64 --ctx 8192 > gpurun_out/step_l70_tp$N.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/step_l70_tp$N.log
done
echo "== bench N=8"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 8 --steps 30 --warmup 3 > gpurun_out/bench_n8.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/bench_n8.log | cut -c1-300
echo "== long ctx tp8"; timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29542 tools/bench_decode_step.py --shape llama3-8b --layers 4 --batch 8 --ctx 131072 > gpurun_out/step_l8_long_tp8.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/step_l8_long_tp8.log
