#!/bin/bash
# Session-13 call 3 (1 GPU): ncu --set full of the prefill attention kernel (what bounds a KV tile?).
mkdir -p gpurun_out
timeout 300 python tools/profile_kernels.py prefill > gpurun_out/prof_prefill_plain.log 2>&1; echo "plain rc=$?"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:prefill_attn -s 1 -c 1 -f -o gpurun_out/prefill_r01c python tools/profile_kernels.py prefill > gpurun_out/ncu_prefill.log 2>&1; echo "rc=$?"; tail -3 gpurun_out/ncu_prefill.log
ls -la gpurun_out/*.ncu-rep
