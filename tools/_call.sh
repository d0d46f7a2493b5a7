mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on -k "regex:greedy_argmax|paged_decode|prefill_|stream_gemm|topk_topp" -o gpurun_out/round2 -f python tools/profile_round2.py > gpurun_out/ncu_round2.log 2>&1; echo "ncu rc=$?"; tail -2 gpurun_out/ncu_round2.log; ls -la gpurun_out/round2.ncu-rep
