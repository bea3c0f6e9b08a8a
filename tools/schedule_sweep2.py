import os, sys, torch
sys.path.insert(0, "/root/repo")
import binius_ntt_b200 as b2
def t(field, log_h, mode):
    if mode is None: os.environ.pop("B2NTT_UPPER", None)
    else: os.environ["B2NTT_UPPER"] = mode
    plan = b2.Plan(field, log_h, 0)
    x = torch.randint(-2**31, 2**31 - 1, (plan.in_words,), dtype=torch.int32, device="cuda")
    for _ in range(3): plan.exec(x, x)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ts = []
    for _ in range(9):
        e0.record(); plan.exec(x, x); e1.record(); torch.cuda.synchronize(); ts.append(e0.elapsed_time(e1))
    plan.close()
    return sorted(ts)[4]
for field, logs in ((128, (21, 22, 23, 24, 25)), (32, (23, 24, 25, 26, 27))):
    for L in logs:
        print(field, L, "auto %.3f  bundle %.3f  tile %.3f" % (t(field, L, None), t(field, L, "0"), t(field, L, "2")), flush=True)
