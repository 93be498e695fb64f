#!/bin/bash
# compute-sanitizer pass over the hand-rolled synchronisation (mbarrier protocols, aliased staging / raw rings, shared
# workspace pool): the smoke decode (both modes, B=2 x T=12) and the tcgen05 kernel tests at small shapes.
#   usage (ONE tool per gpurun call - B200_PROFILING.md): experiments/sanitize.sh memcheck|racecheck|initcheck|synccheck
# Output: gpurun_out/sanitizer_<tool>.txt (summarised in profiles/r02_sanitizer.txt).
TOOL=${1:-memcheck}
cd "$(dirname "$0")/.."
OUT=gpurun_out/sanitizer_${TOOL}.txt
mkdir -p gpurun_out
{
  echo "== compute-sanitizer --tool $TOOL: smoke()"
  timeout 900 compute-sanitizer --tool "$TOOL" --error-exitcode 99 --print-limit 20 python __graft_entry__.py --smoke 2>&1 | tail -25
  echo "smoke rc=${PIPESTATUS[0]}"
  echo "== compute-sanitizer --tool $TOOL: fused / persistent tcgen05 conv kernel tests"
  timeout 1500 compute-sanitizer --tool "$TOOL" --error-exitcode 99 --print-limit 20 \
    python -m pytest tests/test_gpu_kernels.py -q -x -m gpu -k "tc or fused or adain" 2>&1 | tail -25
  echo "kernel tests rc=${PIPESTATUS[0]}"
} > "$OUT" 2>&1
tail -40 "$OUT"
