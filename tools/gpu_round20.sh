#!/bin/bash
# Session-13 call 8 (1 GPU): stagger sweep of the pair kernel.
mkdir -p gpurun_out
for d in 0 800 1500 2200; do
echo "== stagger $d"; B200_PREFILL_STAGGER_CLK=$d timeout 300 python tools/bench_prefill.py l8_b1_s8192,l8_b4_s2048,q7_b2_s4096 2>&1 | grep -oE "'case': '[a-z0-9_]+'|'sm_poly25_tflops': [0-9.]+|'pair2p_poly25_tflops': [0-9.]+|'pair_packed_tflops': [0-9.]+" | paste - - - -
B200_PREFILL_STAGGER_CLK=$d timeout 300 python tools/trace_prefill.py 9 8192 2>/dev/null > gpurun_out/trace_st$d.log; head -1 gpurun_out/trace_st$d.log
done > gpurun_out/stagger.log 2>&1
