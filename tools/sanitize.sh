#!/bin/bash
# compute-sanitizer over every kernel family at small sizes: memcheck and racecheck on tools/sanitize_target.py.
# The kernels synchronise many phases with __syncwarp only (warp-owned lines); racecheck is the tool that would see a hole.
# usage (GPU box): tools/sanitize.sh <outdir>
OUT=${1:-gpurun_out/sanitize}
mkdir -p "$OUT"
cd "$(dirname "$0")/.."
for tool in memcheck racecheck; do
  timeout 1500 compute-sanitizer --tool $tool --print-limit 20 --log-file "$OUT/$tool.log" python tools/sanitize_target.py > "$OUT/$tool.out" 2>&1
  echo "$tool exit $?" >> "$OUT/$tool.out"
  tail -3 "$OUT/$tool.log"
done
