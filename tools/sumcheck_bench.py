"""Sumcheck prover timing on one B200: the whole protocol (num_vars rounds of this_round_messages +
move_to_next_round, then the final claim) on device-resident compact input, the way sumcheck/bench/benchmark.cu times its
"raw" phase, next to the reference's own CUDA prover (oracle/_ref/ref_sumcheck, when built).  One JSON line per case."""
import json, os, subprocess, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
import binius_ntt_b200 as b2


def ours(nv, comp, reps=3):
    words = 4 * comp << nv
    src = torch.randint(-2**31, 2**31 - 1, (words,), dtype=torch.int32, device="cuda")
    rng = np.random.default_rng(1)
    chal = rng.integers(0, 2**32, (nv, 4), dtype=np.uint32)
    best, best_t = None, None
    for _ in range(reps + 1):
        d = src.clone()
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        sc = b2.Sumcheck(d, nv, comp, data_is_transposed=False)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        for r in range(nv):
            sc.this_round_messages()
            sc.move_to_next_round(chal[r])
        sc.this_round_messages()
        t2 = time.perf_counter()
        launches = sc.launches
        sc.close()
        if best is None or t2 - t1 < best:
            best, best_t = t2 - t1, t1 - t0
    # field products of the protocol: per pair of evaluations (C + 1)(C - 1) for the round polynomial + C for the fold
    pairs = (1 << nv) - 1
    prods = pairs * ((comp + 1) * (comp - 1) + comp)
    return {"impl": "b2ntt_sumcheck", "num_vars": nv, "composition_size": comp, "transpose_ms": round(best_t * 1e3, 3), "protocol_ms": round(best * 1e3, 3),
            "gf128_products": prods, "products_per_s": prods / best, "launches": int(launches)}


def reference(nv, comp):
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_sumcheck")
    if not os.path.exists(exe):
        return None
    try:
        out = subprocess.run([exe, str(nv), str(comp), "quiet"], check=True, capture_output=True, text=True, timeout=900).stdout
        return json.loads([ln for ln in out.splitlines() if ln.startswith("{")][-1])
    except Exception as e:  # a slow or failing baseline must not hide our own line
        return {"impl": "reference CUDA sumcheck", "num_vars": nv, "composition_size": comp, "error": str(e)[:200]}


if __name__ == "__main__":
    cases = [(20, 2), (20, 3), (20, 4), (24, 2), (24, 3), (24, 4), (28, 2), (28, 3), (28, 4)]
    if len(sys.argv) > 1:
        cases = [tuple(int(v) for v in a.split(",")) for a in sys.argv[1:]]
    ref_cases = set(c for c in cases if c[0] <= 24 or c == (28, 2))
    for nv, comp in cases:
        print(json.dumps(ours(nv, comp)), flush=True)
        if (nv, comp) in ref_cases:
            r = reference(nv, comp)
            if r:
                print(json.dumps(r), flush=True)
