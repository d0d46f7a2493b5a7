#!/usr/bin/env bash
# Race / memory checks of the hand-written kernels on small shapes (SURVEY.md §5 "race detection").
# GPU box only:  gpurun --timeout 1500 -- 'bash tools/sanitize.sh'   (round 3: the B200 pool refuses compute-sanitizer
# outright, every leg returned rc=86 with the pool's notice)
# Each leg runs under its own `timeout`; logs and a one-line-per-leg summary land in gpurun_out/sanitizer/.
# What is covered: smoke() (split-KV decode with the cluster/DSMEM merge, tcgen05 prefill, RMSNorm, SwiGLU, arg-max,
# split-K GEMM) and a selection of the smallest parity tests of every file (balanced stream-K decode, store_kvcache,
# top-k/top-p, arg-max with library-resident scratch, stream-K tcgen05 GEMM, emulated-rank all-reduce forms on one GPU).
# memcheck = out-of-bounds / misaligned global, shared and local accesses; racecheck = shared-memory hazards inside a
# CTA (it does not see global-memory protocols between CTAs: those are covered by the repeated-replay parity tests);
# synccheck = invalid barrier usage.
set -u
cd "$(dirname "$0")/.."
OUT=gpurun_out/sanitizer
mkdir -p "$OUT"
CS=${CS:-/usr/local/cuda/bin/compute-sanitizer}
LEG_TIMEOUT=${LEG_TIMEOUT:-420}
SMOKE='import __graft_entry__ as g; g.smoke()'
SMALL_TESTS=(
  "tests/test_zz_no_alloc.py::test_no_alloc_matches_workspace_entry_twice_in_a_row"
  "tests/test_gpu_elementwise.py"
  "tests/test_gpu_sampling.py"
)
K_SMALL='not full and not 128k and not bench and not 8192 and not 16384 and not 1500 and not 1000'
: > "$OUT/summary.txt"
leg() {  # name, tool, command...
  local name=$1 tool=$2; shift 2
  local log="$OUT/${name}_${tool}.log"
  timeout "$LEG_TIMEOUT" "$CS" --tool "$tool" --error-exitcode 99 --print-limit 20 "$@" > "$log" 2>&1
  local rc=$?
  local errs
  errs=$(grep -E "ERROR SUMMARY|RACECHECK SUMMARY" "$log" | tail -1)
  echo "$name $tool rc=$rc ${errs:-no summary line (timeout or tool unsupported)}" | tee -a "$OUT/summary.txt"
}
for tool in memcheck racecheck synccheck; do
  leg smoke "$tool" python -c "$SMOKE"
done
leg small_tests memcheck python -m pytest -x -q -m gpu -k "$K_SMALL" "${SMALL_TESTS[@]}"
leg decode_small memcheck python -m pytest -x -q -m gpu tests/test_gpu_decode.py -k "class_api or lse_combine or store_kvcache"
leg decode_small racecheck python -m pytest -x -q -m gpu tests/test_gpu_decode.py -k "class_api or lse_combine"
leg allreduce_emulated memcheck python -m pytest -x -q -m gpu tests/test_gpu_allreduce_norm.py -k "emulated and 896"
leg allreduce_emulated racecheck python -m pytest -x -q -m gpu tests/test_gpu_allreduce_norm.py -k "emulated and 896"
leg stream_gemm memcheck python -m pytest -x -q -m gpu tests/test_gpu_stream_gemm.py -k "grid_sizes or 20-512-64 or 64-256-8"
cat "$OUT/summary.txt"
