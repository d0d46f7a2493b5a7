mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1"
timeout 400 $TR --master-port 29541 bench.py --gpus 8 --steps 20 --warmup 3 > gpurun_out/bench_n8.log 2>&1; echo "bench n8 rc=$?"; grep '^{' gpurun_out/bench_n8.log | cut -c1-3300
timeout 200 $TR --master-port 29542 tools/bench_decode_step.py --profile > gpurun_out/step_l70_tp8_profile.log 2>&1; echo "profile rc=$?"; grep -E '^\{|us/layer' gpurun_out/step_l70_tp8_profile.log | cut -c1-700
B200_AR_MODE=nvlsp timeout 150 $TR --master-port 29543 tools/bench_decode_step.py > gpurun_out/step_l70_tp8_nvlsp.log 2>&1; echo "nvlsp rc=$?"; grep '^{' gpurun_out/step_l70_tp8_nvlsp.log | cut -c1-400
B200_AR_MODE=nvls1 timeout 150 $TR --master-port 29544 tools/bench_decode_step.py > gpurun_out/step_l70_tp8_nvls1.log 2>&1; echo "nvls1 rc=$?"; grep '^{' gpurun_out/step_l70_tp8_nvls1.log | cut -c1-400
timeout 150 $TR --master-port 29545 tools/bench_allreduce.py 64 > gpurun_out/allreduce_tp8.log 2>&1; echo "ar rc=$?"; grep '^{' gpurun_out/allreduce_tp8.log | python -c "import sys,json; j=json.loads(sys.stdin.read()); print({k:v for k,v in j.items() if k.endswith('_us')})"
