#!/bin/bash
# "Reference CUDA" baseline on the B200 box: the reference's own additive_ntt_kernel (GF(2^32) only, the one
# instantiation that exists), built by oracle/Makefile into oracle/_ref/ref_antt from the sources where they lie.
# Prints one JSON line per shape; tools/quick_time.py prints ours for the same shapes.
set -e
cd "$(dirname "$0")/.."
for shape in "20 0" "24 0" "28 0" "20 2" "24 2"; do
  ./oracle/_ref/ref_antt $shape 5
done
python - <<'PY'
import sys; sys.path.insert(0, '.')
from tools.quick_time import time_plan
for lh, lr in [(20, 0), (24, 0), (28, 0), (20, 2), (24, 2)]:
    time_plan(32, lh, lr)
PY
