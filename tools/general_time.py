"""General-basis GF(2^128) additive NTT (b2ntt_general_exec): device-resident time per transform on a random
basis, next to the GF(2^32)-twiddle fast path (b2ntt_exec) on the same size.  Algorithmic traffic = 32 B per element."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np
import torch
import binius_ntt_b200 as b2


def timed(fn, iters):
    for _ in range(2):
        fn()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        fn()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / iters


def run(log_h, iters=3):
    rng = np.random.default_rng(log_h)
    basis = rng.integers(0, 2**32, size=(log_h, 4), dtype=np.uint32)  # independent with probability ~1 - 2^(log_h-128)
    g, plan = b2.GeneralNTT(log_h, basis), b2.Plan(128, log_h, 0)
    x = torch.randint(-2**31, 2**31 - 1, (4 << log_h,), dtype=torch.int32, device="cuda")
    ms_g = timed(lambda: g.exec(x), iters)
    ms_f = timed(lambda: plan.exec(x), iters)
    n = 1 << log_h
    print(json.dumps({"config": "general-basis GF(2^128) additive NTT", "log_h": log_h, "launches": (log_h + 1) // 2 if log_h >= 11 else log_h, "ms": round(ms_g, 3),
                      "elements_per_s": n / ms_g * 1e3, "alg_GBps": round(32 * n / ms_g / 1e6, 1),
                      "fast_path_ms_same_size": round(ms_f, 3)}), flush=True)
    g.close(), plan.close()


if __name__ == "__main__":
    for log_h in ([int(a) for a in sys.argv[1:]] or [16, 20, 24]):
        run(log_h)
