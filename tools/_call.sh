mkdir -p gpurun_out
nvidia-smi -L | head -2
python tests/golden/make_golden_gpu.py > gpurun_out/golden_gpu.log 2>&1; echo "golden rc=$?"; tail -12 gpurun_out/golden_gpu.log
cp gpurun_out/flash_attn_gpu.pt tests/golden/flash_attn_gpu.pt
bash tools/run_gpu_suite.sh
