"""Config 5: bit-sliced GF(2^128) tower multiply microbench, 2^log_n element pairs (default 2^28 = 3 x 4 GiB), plus the
compact and GF(2^32) variants.  48 B of algorithmic traffic per GF(2^128) product (two reads, one write)."""
import json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import binius_ntt_b200 as b2


def run(field_bits, layout, log_n, iters=5):
    n = 1 << log_n
    words = n * field_bits // 32
    a = torch.randint(-2**31, 2**31 - 1, (words,), dtype=torch.int32, device="cuda")
    b = torch.randint(-2**31, 2**31 - 1, (words,), dtype=torch.int32, device="cuda")
    o = torch.empty_like(a)
    for _ in range(2):
        b2.mul(a, b, field_bits, layout, out=o)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(iters):
        b2.mul(a, b, field_bits, layout, out=o)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / iters
    gbs = 3 * words * 4 / ms / 1e6
    print(json.dumps({"config": "tower multiply", "field_bits": field_bits, "layout": "bitsliced" if layout else "compact", "log_n": log_n,
                      "ms": round(ms, 4), "products_per_s": n / ms * 1e3, "alg_GBps": round(gbs, 1), "hbm_frac_of_6447.8": round(gbs / 6447.8, 4)}), flush=True)


if __name__ == "__main__":
    log_n = int(sys.argv[1]) if len(sys.argv) > 1 else 28
    run(128, 1, log_n)
    run(128, 0, log_n)
    run(32, 1, log_n)
    run(32, 0, log_n)
