mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_allreduce_norm.py tests/test_gpu_sampling.py tests/test_gpu_model.py -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
W=2
run() { name=$1; shift; env "$@" timeout 240 python -m torch.distributed.run --nproc-per-node $W --master-addr 127.0.0.1 --master-port 29575 tools/bench_decode_step.py --shape llama3-70b --layers 16 --batch 64 --ctx 8192 $EXTRA > gpurun_out/step_l70_tp${W}_$name.log 2>&1; echo "step $name rc=$?"; grep '^{' gpurun_out/step_l70_tp${W}_$name.log | python -c "import sys,json; [print({k:d.get(k) for k in ('ms_per_step_run','head_ms','tokens_per_s_run','allreduce','lm_head')}) for d in map(json.loads, sys.stdin)]"; }
run headshard
run headrepl B200_HEAD_SHARD=0
timeout 300 python -m torch.distributed.run --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29577 tools/check_tp.py > gpurun_out/check_tp2.log 2>&1; echo "check_tp rc=$?"; tail -3 gpurun_out/check_tp2.log | cut -c1-600
for s in 4 5 6; do timeout 120 python tools/trace_decode.py --shape l70tp8 --splits $s >> gpurun_out/trace_decode.log 2>&1; done; timeout 120 python tools/trace_decode.py --shape l70tp8 --splits 4 --variant 1 >> gpurun_out/trace_decode.log 2>&1; timeout 120 python tools/trace_decode.py --shape l8 >> gpurun_out/trace_decode.log 2>&1; grep '^{' gpurun_out/trace_decode.log | cut -c1-1800
for v in "a" "b B200_STREAM_GEMM=0" "c B200_STREAM_GEMM=1" "d B200_STREAM_GEMM=1 B200_DECODE_PDL=1" "e B200_DECODE_PDL=1" "f B200_DECODE_SPLITS=5" "g B200_DECODE_SPLITS=6" "h B200_DECODE_VARIANT=1"; do set -- $v; tag=$1; shift; env "$@" timeout 200 python tools/profile_layer.py --shape llama3-70b --tp 8 --layers 16 --no-profile --tag "$tag $*" 2>&1 | grep '^{' ; done
