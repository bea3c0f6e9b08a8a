#!/bin/bash
# compare register-budget / CTA-size variants of the pass kernels (development aid)
for W in 8 12; do
 for T in 64 128 256; do
  if [ $W = 8 ]; then unset B2NTT_LIB; else export B2NTT_LIB=$PWD/tools/variants/libb2ntt_w$W.so; fi
  echo "== W=$W T=$T G=${G:-512}"
  B2NTT_CTA_THREADS=$T B2NTT_GROUP_LINES=${G:-512} python - <<'PY'
import sys; sys.path.insert(0, '.')
from tools.quick_time import time_plan
for fb, lh in [(32, 28), (128, 24), (128, 28)]:
    time_plan(fb, lh)
PY
 done
done
