"""Ad-hoc device timing of b2ntt_exec for a few shapes (development aid; bench.py is the contract)."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import binius_ntt_b200 as b2

def time_plan(field_bits, log_h, log_rate=0, layout=0, batch=1, iters=5):
    plan = b2.Plan(field_bits, log_h, log_rate, layout=layout, batch=batch)
    x = torch.randint(-2**31, 2**31 - 1, (plan.in_words,), dtype=torch.int32, device="cuda")
    y = x if log_rate == 0 else torch.empty(plan.out_words, dtype=torch.int32, device="cuda")
    for _ in range(2):
        plan.exec(x, y)
    torch.cuda.synchronize()
    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record()
    for i in range(iters):
        plan.exec(x, y)
        ev[i + 1].record()
    torch.cuda.synchronize()
    ts = [ev[i].elapsed_time(ev[i + 1]) for i in range(iters)]
    ms = sorted(ts)[len(ts) // 2]
    n_out = batch << (log_h + log_rate)
    bytes_alg = (plan.in_words + plan.out_words) * 4
    print(json.dumps(dict(field=field_bits, log_h=log_h, log_rate=log_rate, layout=layout, batch=batch, passes=plan.passes,
                          ms=round(ms, 4), gel_s=round(n_out / ms / 1e6, 3), alg_gbs=round(bytes_alg / ms / 1e6, 1))), flush=True)
    if os.environ.get("B2_PASSES"):
        print("   passes:", " ".join(f"{n.split('::')[1][:6]}={t:.2f}" for n, t in plan.exec_profile(x, y)), flush=True)
    plan.close()

def main():
    shapes = [(32, 20), (32, 24), (32, 28), (128, 20), (128, 24), (128, 26), (128, 28)]
    for fb, lh in shapes:
        time_plan(fb, lh)
    time_plan(32, 20, 2, batch=16)
    time_plan(128, 24, 0, layout=1)
    logic_peak()


def logic_peak(iters=200, lop3_per_mul=990):
    """Dependent-chain bit-sliced GF(2^32) products in registers: LOP3 issue rate of the ALU pipe."""
    nthreads = 148 * 1024
    io = torch.randint(-2**31, 2**31 - 1, (nthreads * 32,), dtype=torch.int32, device="cuda")
    b2.mul_repeat(io, 32, 10)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    b2.mul_repeat(io, 32, iters)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    ops = nthreads * iters * lop3_per_mul
    print(json.dumps(dict(probe="mul_repeat32", ms=round(ms, 3), lop3_per_s=ops / ms * 1e3,
                          lop3_per_clk_per_sm_at_1965MHz=ops / ms * 1e3 / 148 / 1.965e9)), flush=True)


if __name__ == "__main__":
    if len(sys.argv) >= 3:  # quick_time.py FIELD LOG_H [ITERS]: one shape (for ncu launch lists)
        time_plan(int(sys.argv[1]), int(sys.argv[2]), iters=int(sys.argv[3]) if len(sys.argv) > 3 else 2)
    else:
        main()
