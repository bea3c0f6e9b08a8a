#!/bin/bash
# timing sweep over the CTA size knob (development aid)
for T in 64 128 256; do
  echo "== T=$T"
  B2NTT_CTA_THREADS=$T python - <<'PY'
import sys; sys.path.insert(0, '.')
from tools.quick_time import time_plan
for fb, lh in [(32, 24), (32, 28), (128, 24), (128, 28)]:
    time_plan(fb, lh)
PY
done
