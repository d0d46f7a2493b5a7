#!/bin/bash
# Session-5 call 5 (8 GPUs): fused peer-memory all-reduce + norm at TP=8 / TP=4: one-shot vs row-sharded vs NCCL.
mkdir -p gpurun_out
run() { # name nproc port env... -- args
  name=$1; np=$2; port=$3; shift 3
  echo "== $name"; env "$@" timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node $np --master-addr 127.0.0.1 --master-port $port tools/bench_decode_step.py $ARGS > gpurun_out/$name.log 2>&1; echo "rc=$?"; tail -1 gpurun_out/$name.log | cut -c1-400
}
echo "== peer pytest (2 procs)"; timeout 300 python -m pytest tests/test_gpu_allreduce_norm.py -m gpu -q --timeout 280 -p no:cacheprovider -k "peer_memory" > gpurun_out/pytest_peer.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/pytest_peer.log | cut -c1-300
echo "== check_tp N=8 sharded=1"; B200_AR_SHARDED=1 timeout 300 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29601 tools/check_tp.py > gpurun_out/check_tp8_sh1.log 2>&1; echo "rc=$?"; grep -E "^\{" gpurun_out/check_tp8_sh1.log
ARGS="--shape llama3-70b --layers 16 --batch 64 --ctx 8192"
run step_l70_tp8_sharded 8 29611 B200_AR_SHARDED=1
run step_l70_tp8_oneshot 8 29612 B200_AR_SHARDED=0
run step_l70_tp8_nccl 8 29613 B200_FUSED_ALLREDUCE=0
run step_l70_tp4_sharded 4 29614 B200_AR_SHARDED=1
run step_l70_tp4_oneshot 4 29615 B200_AR_SHARDED=0
ARGS="--shape llama3-8b --layers 4 --batch 8 --ctx 131072"
run step_l8_long_tp8_sharded 8 29616 B200_AR_SHARDED=1
