"""Run a few device-resident transforms of one shape (profiling target for ncu):
    python tools/one_transform.py [log_h=26] [field_bits=128] [count=2]"""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import binius_ntt_b200 as b2

log_h = int(sys.argv[1]) if len(sys.argv) > 1 else 26
field = int(sys.argv[2]) if len(sys.argv) > 2 else 128
count = int(sys.argv[3]) if len(sys.argv) > 3 else 2
plan = b2.Plan(field, log_h, 0)
x = torch.randint(-2**31, 2**31 - 1, (plan.in_words,), dtype=torch.int32, device="cuda")
for _ in range(count):
    plan.exec(x, x)
torch.cuda.synchronize()
print("passes:", " ".join(f"{n.split('::')[1]}={t:.3f}" for n, t in plan.exec_profile(x, x)))
