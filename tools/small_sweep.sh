# development aid: the two pass schedules (B2NTT_UPPER=2: upper + tile passes, 0: bundle passes only) across sizes
python - <<'PY'
import os, sys, subprocess, json
sys.path.insert(0, os.getcwd())
shapes = [(32, 20, 2, 32), (32, 20, 2, 256), (32, 24, 0, 1), (32, 26, 0, 1), (32, 28, 0, 1), (128, 23, 0, 1), (128, 25, 0, 1), (128, 26, 0, 1), (32, 16, 0, 64), (128, 20, 0, 16)]
code = """
import sys, os
sys.path.insert(0, os.getcwd())
sys.path.insert(0, os.path.join(os.getcwd(), 'tools'))
import quick_time as q
q.time_plan(%d, %d, %d, batch=%d, iters=7)
"""
for s in shapes:
    for u in ("2", "0"):
        env = dict(os.environ, B2NTT_UPPER=u)
        out = subprocess.run([sys.executable, "-c", code % s], env=env, capture_output=True, text=True).stdout.strip().splitlines()
        print("UPPER=" + u, out[0] if out else "failed", flush=True)
PY
