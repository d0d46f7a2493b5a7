mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_decode.py tests/test_gpu_model.py tests/test_gpu_vs_flash_attn.py -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
rm -f gpurun_out/trace_decode.log
for s in 4; do timeout 120 python tools/trace_decode.py --shape l70tp8 --splits $s >> gpurun_out/trace_decode.log 2>&1; B200_DECODE_NO_CLUSTER=1 timeout 120 python tools/trace_decode.py --shape l70tp8 --splits $s >> gpurun_out/trace_decode.log 2>&1; done; grep '^{' gpurun_out/trace_decode.log | cut -c1-1700
for v in "a" "b B200_DECODE_NO_CLUSTER=1" "c B200_DECODE_SPLITS=5" "d B200_DECODE_SPLITS=6" "e B200_DECODE_PDL=1" "f B200_DECODE_PDL=2"; do set -- $v; tag=$1; shift; env "$@" timeout 200 python tools/profile_layer.py --shape llama3-70b --tp 8 --layers 16 --no-profile --tag "$tag $*" 2>&1 | grep '^{' | cut -c1-200; done
timeout 200 python tools/tune_decode.py --shape l70tp8 --variants 6,1 --splits 2,3,4,5,6,8 --no-fa 2>&1 | tail -3 | cut -c1-1500
