#!/bin/bash
# Session-13 call 4 (1 GPU): in-kernel pipeline trace of the prefill kernel (scalar and packed softmax).
mkdir -p gpurun_out
for v in 2 3 5; do timeout 300 python tools/trace_prefill.py $v 8192 > gpurun_out/trace_v$v.log 2>&1; echo "v$v rc=$?"; head -1 gpurun_out/trace_v$v.log; done
