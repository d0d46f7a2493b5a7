#!/bin/bash
# Stage the reference's three Triton kernel files, unmodified, where the GPU box can import them
# (baseline/_ref/ is git-ignored but travels with the gpurun snapshot).  Build-container only:
# /root/reference does not exist on the GPU box.  Used by tests/golden/make_golden_triton.py.
set -e
ROOT="$(cd "$(dirname "$0")/../.." && pwd)"
REF="${REFERENCE_ROOT:-/root/reference}"
mkdir -p "$ROOT/baseline/_ref/kernel"
cp "$REF/kernel/swiglu.py" "$REF/kernel/rmsnorm_add.py" "$REF/kernel/rmsnorm_residual.py" "$ROOT/baseline/_ref/kernel/"
ls -la "$ROOT/baseline/_ref/kernel"
