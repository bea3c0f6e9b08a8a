#!/usr/bin/env python3
"""bench.py -- headline benchmark: GF(2^128) additive-NTT elements/s at log_n = 28 on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--log-n L]

A step = one forward additive NTT (b2ntt_exec) of 2^log_n GF(2^128) elements, compact layout,
in place, input resident in HBM.  The data set (4 GiB at log_n = 28) is far larger than the 126 MB
L2, so no flush is needed between steps.  At N > 1 (torchrun, one rank per GPU) every rank runs an
independent replica of the same transform ("replicas only", DESIGN.md section e; weak scaling);
value = N * 2^log_n * K / max-over-ranks device time.

Beside the headline the line carries a `configs` block: the other BASELINE.json configurations, each
timed on the device and VERIFIED (golden digests of the reference's own test table, known-answer
products, block-by-block parity of the split transform) -- config 1, 2 and 5 on rank 0 at N = 1;
config 3 (256 RS columns sharded over the ranks, no collective) at every N; config 4 (one transform
split four-step over the ranks, one all-to-all over NVLink) at N >= 2.

`--impl reference` times the CPU path instead: the reference has no CPU additive NTT, so this is the
oracle port (oracle/antt.c, all host threads) on a bounded sample of the same workload.
"""
import argparse
import hashlib
import json
import math
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

METRIC = "GF(2^128) additive-NTT elements/s"
UNIT = "elements/s"
BYTES_PER_ELEMENT = 32  # one read + one write of a 16-byte element per transform
KERNEL_KEYS = {"b2::bottom_kernel": "bottom_kernel", "b2::upper_kernel": "upper_kernel", "b2::pass_kernel": "pass_kernel"}


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        return {"hbm_gbs": 6550.0, "sm_max_mhz": 1965.0}, "fallback of B200_PROFILING.md (6550 GB/s)"


def host_threads():
    """Threads the CPU arm may use: the affinity mask, NOT OMP_NUM_THREADS (torchrun sets that to 1)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def golden_md5(log_rate, log_h):
    """The reference's digest table (ntt/tests/test_ntt.cu:52-124) as committed data: tests/golden/additive_ntt_md5.json."""
    with open(os.path.join(ROOT, "tests", "golden", "additive_ntt_md5.json")) as f:
        return json.load(f)["md5"][str(log_rate)].get(str(log_h))


def reference_stream(log_h, log_rate):
    """Input of ntt/tests/test_ntt.cu:192-199: std::mt19937(0xdeadbeef + log_h + log_rate), one draw per element
    (numpy's legacy RandomState seeds and draws exactly like std::mt19937)."""
    import numpy as np

    return np.random.RandomState((0xDEADBEEF + log_h + log_rate) & 0xFFFFFFFF).randint(0, 2**32, 1 << log_h, dtype=np.uint32)


def md5_words(t):
    """MD5 over the little-endian words of a tensor / array, as the reference's test hashes its output (test_ntt.cu:208-214)."""
    import numpy as np

    a = t.cpu().numpy() if hasattr(t, "cpu") else t
    return hashlib.md5(np.ascontiguousarray(a).view(np.uint8)).hexdigest()


class ClockSampler(threading.Thread):
    """Samples SM clocks and throttle reasons with nvidia-smi while the timed region runs."""

    def __init__(self, gpu_index):
        super().__init__(daemon=True)
        self.gpu, self.samples, self.reasons, self.max_mhz, self._halt = gpu_index, [], set(), None, threading.Event()

    def run(self):
        q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
            "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        while not self._halt.is_set():
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={q}", "--format=csv,noheader,nounits", "-i", str(self.gpu)],
                                     capture_output=True, text=True, timeout=5).stdout.strip().split(",")
                self.samples.append(float(out[0]))
                self.max_mhz = float(out[1])
                for n, v in zip(names, out[2:]):
                    if v.strip().lower() == "active":
                        self.reasons.add(n)
            except Exception:
                pass
            self._halt.wait(0.05)

    def finish(self):
        self._halt.set()
        self.join(timeout=6)
        s = sorted(self.samples)
        return {"sm_mhz": s[len(s) // 2] if s else None, "sm_max_mhz": self.max_mhz, "reasons": sorted(self.reasons),
                "samples": len(s)}


def config_dict(args, extra=None):
    d = {"workload": f"additive NTT log_n={args.log_n} GF(2^128) compact layout, log_rate=0, one column per GPU, in place",
         "log_n": args.log_n, "field": "GF(2^128) tower, 4 x u32 limbs", "l2_policy": "data set (4 GiB at log_n=28) >> 126 MB L2; no flush",
         "parallelism": "replicas only (one independent transform per GPU)"}
    if extra:
        d.update(extra)
    return d


def cpu_sample(log_h, repeats=1):
    """Time the CPU oracle port (the checker, used here only as the timed CPU baseline) on a bounded
    sample: a GF(2^128) forward transform of 2^log_h elements on all host threads."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import oracle as O

    threads = host_threads()
    x = np.random.default_rng(1).integers(0, 2**32, size=(1 << log_h) * 4, dtype=np.uint32)
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        O.antt_forward(x, log_h, 0, 4, threads)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return {"value": (1 << log_h) / best, "unit": UNIT, "cores": threads, "kind": "port",
            "sample": f"one GF(2^128) additive NTT, log_n={log_h}, oracle/antt.c with OpenMP on {threads} host threads "
                      f"({best:.2f} s); the reference has no CPU NTT (SURVEY F5)"}


def reference_cuda(log_h=28):
    """The reference's own CUDA kernels on this GPU (oracle/_ref/ref_antt = ntt/additive_ntt.cuh with a plain main, built by
    oracle/Makefile): GF(2^32) only -- the one instantiation the reference has.  A baseline, never the product."""
    exe = os.path.join(ROOT, "oracle", "_ref", "ref_antt")
    if not os.path.exists(exe):
        return {"unavailable": "oracle/_ref/ref_antt was not built (reference sources absent at build time)"}
    try:
        out = subprocess.run([exe, str(log_h), "0", "3"], capture_output=True, text=True, timeout=300).stdout.strip().splitlines()
        d = json.loads(out[-1])
        n = 1 << log_h
        return {"workload": f"additive NTT log_n={log_h} GF(2^32) (the reference's only instantiation), reference kernels on this GPU",
                "kernels_only_ms": d["kernels_ms"], "whole_apply_ms": d["apply_ms"], "kernel_launches": d["kernels"],
                "kernels_only_elements_per_s": n / d["kernels_ms"] * 1e3, "whole_apply_elements_per_s": n / d["apply_ms"] * 1e3,
                "md5_matches_reference_table": d["md5"] == golden_md5(0, log_h)}
    except Exception as e:  # noqa: BLE001
        return {"unavailable": f"ref_antt failed: {e}"}


def run_reference(args, rank, world):
    """--impl reference: the CPU path.  Rank 0 alone works; a step = one bounded sample transform."""
    if rank != 0:
        return
    log_s = min(args.log_n, args.ref_log_n)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import oracle as O

    threads = host_threads()
    x = np.random.default_rng(1).integers(0, 2**32, size=(1 << log_s) * 4, dtype=np.uint32)
    for _ in range(args.warmup):
        O.antt_forward(x, log_s, 0, 4, threads)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        O.antt_forward(x, log_s, 0, 4, threads)
    dt = time.perf_counter() - t0
    val = args.steps * (1 << log_s) / dt
    sample = f"each step = one GF(2^128) additive NTT of 2^{log_s} elements (bounded sample of the log_n={args.log_n} workload), " \
             f"oracle port on {threads} host threads"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32 limbs (GF(2^32) tower arithmetic)", "data": "synthetic",
        "config": config_dict(args, {"cpu_sample_log_n": log_s, "cpu_threads": threads}),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": threads, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }), flush=True)


# ---- helpers of the GPU arm ---------------------------------------------------------------------------------------------------------
class Launches:
    n = 0


def dev_time_ms(fn, iters, stream):
    """Median device time of `fn` over `iters` runs, CUDA events on the launching stream."""
    import torch

    ev = [torch.cuda.Event(enable_timing=True) for _ in range(iters + 1)]
    ev[0].record(stream)
    for i in range(iters):
        fn()
        ev[i + 1].record(stream)
    torch.cuda.synchronize()
    ts = sorted(ev[i].elapsed_time(ev[i + 1]) for i in range(iters))
    return ts[len(ts) // 2]


def bench_small_configs(b2, dev, peak, stream):
    """configs 1, 2 and 5 of BASELINE.json on one GPU: device time + verification."""
    import numpy as np
    import torch

    out = {}
    KATS = json.load(open(os.path.join(ROOT, "tests", "golden", "field_kats.json")))
    # -- config 1: log_n = 20 GF(2^32), one column.  81 us per transform: 64 transforms between two events.
    plan = b2.Plan(32, 20, 0, device=dev.index)
    x = torch.randint(-2**31, 2**31 - 1, (1 << 20,), dtype=torch.int32, device=dev)
    for _ in range(5):
        plan.exec(x, x, stream)
    reps = 64
    ms = dev_time_ms(lambda: [plan.exec(x, x, stream) for _ in range(reps)], 5, stream) / reps
    Launches.n += (5 + 5 * reps) * plan.launches
    got = plan.exec_host(reference_stream(20, 0))
    Launches.n += plan.launches
    out["config1_log20_gf32"] = {"ms": ms, "value": (1 << 20) / ms * 1e3, "unit": UNIT, "roofline_frac": 8 * (1 << 20) / (ms * 1e-3) / 1e9 / peak,
                                 "launches_per_transform": plan.launches, "timing": f"{reps} back-to-back transforms per event pair (L2-resident: 4 MiB)",
                                 "verified": md5_words(got) == golden_md5(0, 20), "check": "MD5 == additive_ntt_hashes[0][20] (ntt/tests/test_ntt.cu:75)"}
    plan.close()
    # -- config 2: log_n = 24 GF(2^128), compact and bit-sliced layouts
    n = 1 << 24
    xin = torch.randint(-2**31, 2**31 - 1, (n * 4,), dtype=torch.int32, device=dev)
    xin.view(-1, 4)[:, 0] = torch.from_numpy(reference_stream(24, 0).view(np.int32)).to(dev)
    for name, layout in (("compact", 0), ("bitsliced", 1)):
        plan = b2.Plan(128, 24, 0, layout=layout, device=dev.index)
        buf = xin.clone()
        if layout:
            b2.bitslice_(buf, 128, stream=stream)
        y = plan.exec(buf, buf, stream)
        if layout:
            b2.bitslice_(y, 128, inverse=True, stream=stream)
        ok = md5_words(y.view(-1, 4)[:, 0].contiguous()) == golden_md5(0, 24)
        for _ in range(3):
            plan.exec(buf, buf, stream)
        ms = dev_time_ms(lambda: plan.exec(buf, buf, stream), 10, stream)
        Launches.n += 14 * plan.launches + (2 if layout else 0)
        out[f"config2_log24_gf128_{name}"] = {"ms": ms, "value": n / ms * 1e3, "unit": UNIT, "roofline_frac": 32 * n / (ms * 1e-3) / 1e9 / peak,
                                              "launches_per_transform": plan.launches, "verified": ok,
                                              "check": "MD5 of limb 0 == additive_ntt_hashes[0][24] (test_ntt.cu:79; limb-wise action, SURVEY F8)"}
        plan.close()
        del buf, y
    del xin
    # -- config 5: 2^28 bit-sliced GF(2^128) products (48 algorithmic bytes per product)
    nb, W = 1 << 23, 128
    a = torch.randint(-2**31, 2**31 - 1, (nb * W,), dtype=torch.int32, device=dev)
    b = torch.randint(-2**31, 2**31 - 1, (nb * W,), dtype=torch.int32, device=dev)
    case = next(c for c in KATS["bitsliced"] if c["h"] == 7 and len(c["a"]) == 16)  # finite_fields/tests/test_fanpaartower.cu:204-273
    ka, kb = np.zeros(W, dtype=np.uint32), np.zeros(W, dtype=np.uint32)
    ka[:16], kb[:16] = case["a"], case["b"]
    for blk in (0, nb - 1):
        for t, k in ((a, ka), (b, kb)):
            seg = t[blk * W:(blk + 1) * W]
            seg.copy_(torch.from_numpy(k.view(np.int32)).to(dev))
            b2.bitslice_(seg, W, stream=stream)
    o = torch.empty_like(a)
    b2.mul(a, b, 128, b2.api.LAYOUT_BITSLICED, out=o, stream=stream)
    ok = True
    for blk in (0, nb - 1):
        seg = o[blk * W:(blk + 1) * W].clone()
        b2.bitslice_(seg, W, inverse=True, stream=stream)
        got = seg.cpu().numpy().view(np.uint32)
        ok = ok and all(int(got[int(i)]) == v for i, v in case["result"].items()) and not got[16:].any()
    # commutativity on the whole buffer: a*b == b*a word for word
    o2 = b2.mul(b, a, 128, b2.api.LAYOUT_BITSLICED, stream=stream)
    ok = ok and bool(torch.equal(o, o2))
    del o2
    ms = dev_time_ms(lambda: b2.mul(a, b, 128, b2.api.LAYOUT_BITSLICED, out=o, stream=stream), 5, stream)
    Launches.n += 7 + 12
    out["config5_mul128_bitsliced_2p28"] = {"ms": ms, "value": (1 << 28) / ms * 1e3, "unit": "products/s", "roofline_frac": 48 * (1 << 28) / (ms * 1e-3) / 1e9 / peak,
                                            "verified": ok, "check": "the four reference products (test_fanpaartower.cu:204-273) planted in the first and last block; a*b == b*a on all 2^28"}
    del a, b, o
    torch.cuda.empty_cache()
    return out


def bench_rs(b2, dev, peak, stream, world, dist, iters=3):
    """config 3: 256 columns x (log_h = 20, rate 4) GF(2^32) sharded by column over the ranks; no collective."""
    import numpy as np
    import torch

    rs = b2.ColumnShardedRS(256, 20, 2, field_bits=32, device=dev.index)
    xin = torch.randint(-2**31, 2**31 - 1, (rs.in_words,), dtype=torch.int32, device=dev)
    xin[:1 << 20] = torch.from_numpy(reference_stream(20, 2).view(np.int32)).to(dev)  # this rank's first column = the reference stream
    xout = torch.empty(rs.out_words, dtype=torch.int32, device=dev)
    rs.encode(xin, xout, stream)
    ok = md5_words(xout[:1 << 22]) == golden_md5(2, 20)
    rs.encode(xin, xout, stream)
    torch.cuda.synchronize()
    if world > 1:
        dist.barrier()
    ms = dev_time_ms(lambda: rs.encode(xin, xout, stream), iters, stream)
    Launches.n += (2 + iters) * rs.plan.launches
    t = torch.tensor([ms, 0.0 if ok else 1.0], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms, ok = float(t[0]), float(t[1]) == 0.0
    res = {"ms": ms, "value": 256 * (1 << 22) / ms * 1e3, "unit": "output symbols/s", "columns_per_gpu": rs.count, "collective": "none",
           "roofline_frac_per_gpu": 20 * rs.count * (1 << 20) / (ms * 1e-3) / 1e9 / peak, "timing": "max over ranks of the median device time",
           "verified": ok, "check": "every rank: MD5 of its first column == additive_ntt_hashes[2][20] = 53e8a086... (test_ntt.cu:112)"}
    rs.close()
    del xin, xout
    torch.cuda.empty_cache()
    return res


def bench_four_step(b2, dev, stream, world, rank, dist, iters=3):
    """config 4: ONE GF(2^128) transform of 2^(27 + log2 N) elements split over the N ranks (top stages, one all-to-all over
    NVLink, local transform), plus block-by-block parity against a single-GPU transform at log_n = 22."""
    import torch

    def max_of(v):
        t = torch.tensor([v], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    log_s = 22
    g = torch.Generator(device=dev).manual_seed(1234)  # the same stream on every rank
    x = torch.randint(-2**31, 2**31 - 1, ((1 << log_s) * 4,), dtype=torch.int32, device=dev, generator=g)
    whole = b2.Plan(128, log_s, 0, device=dev.index)
    ref = whole.exec(x.clone())
    ok = True
    for fused in (True, False):  # both exchange paths against the single-GPU transform
        fs = b2.FourStepNTT(log_s, 128, device=dev.index, fused=fused)
        out = fs.forward(b2.FourStepNTT.m_shard(x, world, rank, 4).contiguous().clone())
        ok = ok and bool(torch.equal(out, ref[rank * fs.local_words:(rank + 1) * fs.local_words]))
        Launches.n += fs.launches
        fs.close()
    Launches.n += whole.launches
    whole.close()
    del x, ref, out
    bad = max_of(0.0 if ok else 1.0)
    log_n = min(30, 27 + int(math.log2(world)))
    # -- A/B baseline: phase 1 in place, NCCL all_to_all_single, phase 2
    fs = b2.FourStepNTT(log_n, 128, device=dev.index, fused=False)
    buf = torch.randint(-2**31, 2**31 - 1, (fs.local_words,), dtype=torch.int32, device=dev)
    scratch = torch.empty_like(buf)
    for _ in range(2):
        fs.forward(buf, scratch)
    torch.cuda.synchronize()
    dist.barrier()
    tot = [0.0, 0.0, 0.0]
    whole_ms = 0.0
    for _ in range(iters):
        e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
        e[0].record(stream)
        fs.top(buf, stream)
        e[1].record(stream)
        fs.exchange(buf, scratch)
        e[2].record(stream)
        fs.local(scratch, stream)
        e[3].record(stream)
        torch.cuda.synchronize()
        for i in range(3):
            tot[i] += e[i].elapsed_time(e[i + 1]) / iters
        whole_ms += e[0].elapsed_time(e[3]) / iters
    Launches.n += (2 + iters) * fs.launches
    ms = [max_of(t) for t in tot]
    nccl_total = max_of(whole_ms)
    bytes_each_way = fs.local_words * 4 * (world - 1) // world
    nccl = {"ms": nccl_total, "ms_top": ms[0], "ms_exchange": ms[1], "ms_local": ms[2], "exchange": fs.exchange_kind,
            "exchange_GBps_per_gpu": bytes_each_way / ms[1] / 1e6}
    fs.close()
    del scratch
    # -- the product path: exchange fused into the phase-1 kernel's peer stores
    fs = b2.FourStepNTT(log_n, 128, device=dev.index, fused=True)
    for _ in range(2):
        fs.forward(buf)
    torch.cuda.synchronize()
    dist.barrier()
    t_sc = t_lo = t_all = 0.0
    for _ in range(iters):
        e = [torch.cuda.Event(enable_timing=True) for _ in range(3)]
        e[0].record(stream)
        fs._stream_barrier()
        fs.scatter(buf)
        fs._stream_barrier()
        e[1].record(stream)
        fs.local(fs.block)
        e[2].record(stream)
        torch.cuda.synchronize()
        t_sc += e[0].elapsed_time(e[1]) / iters
        t_lo += e[1].elapsed_time(e[2]) / iters
        t_all += e[0].elapsed_time(e[2]) / iters
    Launches.n += (2 + iters) * fs.launches
    t_sc, t_lo, total = max_of(t_sc), max_of(t_lo), max_of(t_all)
    res = {"log_n": log_n, "ms": total, "value": (1 << log_n) / total * 1e3, "unit": UNIT, "ms_scatter": t_sc, "ms_local": t_lo,
           "ms_exchange_exposed": max(0.0, t_sc - ms[0]), "exchange": fs.exchange_kind, "exchange_bytes_per_gpu_each_way": bytes_each_way,
           "scatter_GBps_per_gpu": bytes_each_way / t_sc / 1e6, "nccl_baseline": nccl, "timing": "max over ranks, CUDA events per phase; "
           "ms_scatter spans barrier + phase-1 kernel with peer stores + barrier; ms_exchange_exposed = ms_scatter - the baseline's ms_top",
           "verified": bad == 0.0, "check": f"log_n={log_s}, both exchange paths: every rank's block == the same block of a single-GPU transform of the whole input"}
    fs.close()
    del buf
    torch.cuda.empty_cache()
    return res


def logic_roofline(b2, dev, stream, kernels_ms, clocks_mhz):
    """SURVEY 8d's second figure: ALU-pipe (LOP3 / SHF / PRMT) and FMA-pipe (IMAD) work of one transform against the measured
    pipe peak.  Executed instructions per launch come from the ncu capture of this very command (profiles/traffic.json, per
    kernel); the peak is measured in this run with b2ntt_mul_repeat (dependent bit-sliced products in registers: pure LOP3)."""
    import torch

    nthreads, iters, lop3_per_mul = 148 * 1024, 200, 990
    io = torch.randint(-2**31, 2**31 - 1, (nthreads * 32,), dtype=torch.int32, device=dev)
    b2.mul_repeat(io, 32, 10, stream=stream)
    ms = dev_time_ms(lambda: b2.mul_repeat(io, 32, iters, stream=stream), 3, stream)
    Launches.n += 4
    peak_alu = nthreads * iters * lop3_per_mul / (ms * 1e-3)  # lane-ops/s of the ALU pipe
    res = {"peak_alu_lane_ops_per_s": peak_alu, "peak_source": "b2ntt_mul_repeat(32): 990 LOP3 per dependent product, measured in this run",
           "peak_alu_lanes_per_clk_per_sm": peak_alu / 148 / ((clocks_mhz or 1965.0) * 1e6)}
    try:
        with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
            tj = json.load(f)
        alu = fma = total = 0.0
        for name, v in kernels_ms.items():
            k = tj["kernels"][name]
            alu += k["alu_warp_inst_per_launch"] * v["launches_per_transform"]
            fma += k["fma_warp_inst_per_launch"] * v["launches_per_transform"]
            total += k["warp_inst_per_launch"] * v["launches_per_transform"]
        step_s = sum(v["avg_launch_ms"] * v["launches_per_transform"] for v in kernels_ms.values()) * 1e-3
        res.update({"alu_lane_ops_per_step": alu * 32, "fma_lane_ops_per_step": fma * 32, "warp_inst_per_step": total,
                    "frac": alu * 32 / step_s / peak_alu, "fma_frac": fma * 32 / step_s / peak_alu,
                    "issue_frac": total / step_s / (148 * 4 * (clocks_mhz or 1965.0) * 1e6),
                    "counts_source": tj.get("source"), "bound": "alu pipe + issue slots (register-file operand bandwidth caps dual-pipe issue at ~0.76 inst/clk/sub-partition, profiles/README.md)"})
    except Exception as e:  # noqa: BLE001
        res["counts_unavailable"] = str(e)
    return res


def run_ours(args, rank, world, local_rank):
    import numpy as np
    import torch
    import torch.distributed as dist

    import binius_ntt_b200 as b2

    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device; the product has no CPU path")
    b2.lib()  # fail loudly if the CUDA library is missing
    torch.cuda.set_device(local_rank)
    dev = torch.device("cuda", local_rank)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    n = 1 << args.log_n
    plan = b2.Plan(128, args.log_n, 0, device=local_rank)
    launches_per_step = plan.launches
    g = torch.Generator(device=dev).manual_seed(0x5EED + rank)
    x = torch.randint(-2**31, 2**31 - 1, (n * 4,), dtype=torch.int32, device=dev, generator=g)
    stream = torch.cuda.current_stream()
    peaks, peak_src = measured_peaks()
    peak = float(peaks["hbm_gbs"])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    # ---- device-resident throughput -----------------------------------------------------------------
    for _ in range(args.warmup):
        plan.exec(x, x, stream)
    barrier()
    sampler = ClockSampler(local_rank) if rank == 0 else None
    if sampler:
        sampler.start()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(stream)
    for _ in range(args.steps):
        plan.exec(x, x, stream)
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.finish() if sampler else None
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    Launches.n += (args.warmup + args.steps) * launches_per_step
    # per-kernel split of the step: the same launches timed one by one (CUDA events around every pass, b2ntt_exec_profile)
    pass_ms = {}
    prof_steps = min(args.steps, 5)
    for _ in range(prof_steps):
        for name, t_ms in plan.exec_profile(x, x, stream):
            pass_ms.setdefault(name, []).append(t_ms)
    Launches.n += prof_steps * launches_per_step

    # ---- verification of the timed path: the reference's golden digest on limb 0 -----
    verified = None
    if rank == 0 and golden_md5(0, args.log_n):
        x.view(-1, 4)[:, 0] = torch.from_numpy(reference_stream(args.log_n, 0).view(np.int32)).to(dev)
        plan.exec(x, x, stream)
        Launches.n += launches_per_step
        verified = md5_words(x.view(-1, 4)[:, 0].contiguous()) == golden_md5(0, args.log_n)

    # ---- end to end through the reference-facing calls: host in, host out ---------------------------
    # (a) b2ntt_exec_host = AdditiveNTT::apply, one blocking call per step (H2D + kernels + D2H in series);
    # (b) b2ntt_exec_host_batch = the same K steps in one pipelined call (step i copied in while step i-1 is
    #     transformed and step i-2 is copied out).  Every step moves its own 4 GiB in and 4 GiB out of pinned memory.
    e2e_steps = max(1, min(args.steps, 10))
    hin = [torch.empty(n * 4, dtype=torch.int32).pin_memory() for _ in range(2)]
    hout = [torch.empty(n * 4, dtype=torch.int32).pin_memory() for _ in range(2)]
    for hb in hin:
        hb.copy_(x)  # a valid input image
    plan.exec_host(hin[0], hout[0])  # warm-up (allocates the plan's staging buffer)
    barrier()
    t0 = time.perf_counter()
    for i in range(min(e2e_steps, 2)):
        plan.exec_host(hin[i % 2], hout[i % 2])
    torch.cuda.synchronize()
    dt_blocking = (time.perf_counter() - t0) / min(e2e_steps, 2)
    plan.exec_host_batch([hin[0]], [hout[0]])  # warm-up (allocates the pipeline's buffers and streams)
    barrier()
    t0 = time.perf_counter()
    plan.exec_host_batch([hin[i % 2] for i in range(e2e_steps)], [hout[i % 2] for i in range(e2e_steps)])
    torch.cuda.synchronize()
    dt = time.perf_counter() - t0
    t = torch.tensor([dt, dt_blocking], dtype=torch.float64, device=dev)
    tmin = t.clone()
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        dist.all_reduce(tmin, op=dist.ReduceOp.MIN)
    e2e_val = world * n * e2e_steps / float(t[0].item())
    e2e_blocking = world * n / float(t[1].item())
    Launches.n += (min(e2e_steps, 2) + e2e_steps + 2) * launches_per_step
    host_gbs = {"slowest_rank_each_way_GBps": n * 16 * e2e_steps / float(t[0].item()) / 1e9,
                "fastest_rank_each_way_GBps": n * 16 * e2e_steps / float(tmin[0].item()) / 1e9,
                "note": "every rank streams 4 GiB in and 4 GiB out of pinned host memory per step; all GPUs of the box hang off one host "
                        "(one NUMA node, CPU affinity 0-31 for every GPU), so N ranks share the host's memory and PCIe root bandwidth"}
    del hin, hout, x
    plan.close()
    torch.cuda.empty_cache()

    # ---- the other BASELINE.json configurations -------------------------------------------------------------
    configs = {}
    if not args.no_configs:
        if world == 1:
            configs.update(bench_small_configs(b2, dev, peak, stream))
        configs["rs_256col"] = bench_rs(b2, dev, peak, stream, world, dist)
        if world > 1 and (world & (world - 1)) == 0:
            configs["four_step"] = bench_four_step(b2, dev, stream, world, rank, dist)

    if rank == 0:
        # dominant kernel = the one with the largest share of the step; its launches were timed live above
        dom = max(pass_ms, key=lambda k: sum(pass_ms[k]))
        launch_ms = sum(pass_ms[dom]) / len(pass_ms[dom])
        ach = BYTES_PER_ELEMENT * n / (launch_ms * 1e-3) / 1e9     # GB/s, algorithmic bytes of one launch (whole data set in + out)
        step_ms = ms / args.steps
        kernels = {k: {"launches_per_transform": len(v) // prof_steps, "avg_launch_ms": sum(v) / len(v),
                       "share_of_step": sum(v) / prof_steps / step_ms} for k, v in pass_ms.items()}
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                tj = json.load(f)
            if tj.get("log_n") == args.log_n:
                traffic = tj["kernels"][dom]["dram_bytes_per_launch"]
        except Exception:
            pass
        whole = BYTES_PER_ELEMENT * n / (step_ms * 1e-3) / 1e9
        roof = {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s",
                "frac": ach / peak, "traffic": traffic, "peak_source": peak_src, "kernels": kernels,
                "launches_per_transform": launches_per_step, "avg_launch_ms": launch_ms,
                "algorithmic_bytes_per_launch": BYTES_PER_ELEMENT * n,
                "whole_transform_achieved": whole, "whole_transform_frac": whole / peak,
                "limiter": "alu (LOP3) pipe and issue slots, not HBM: see roofline.logic"}
        roof["logic"] = logic_roofline(b2, dev, stream, kernels, clocks["sm_mhz"] if clocks else None)
        cpu = cpu_sample(args.cpu_log_n) if (world == 1 and not args.no_cpu) else None
        line = {
            "metric": METRIC, "value": world * n * args.steps / (ms_max * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32 limbs (GF(2^32) tower arithmetic, bit-sliced in registers)",
            "data": "synthetic", "config": config_dict(args), "verified": verified,
            "verified_check": f"limb 0 fed with the reference stream: MD5 == additive_ntt_hashes[0][{args.log_n}] (ntt/tests/test_ntt.cu:52-124)",
            "roofline": roof,
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": n * 16, "d2h_bytes_per_step": n * 16,
                    "steps": e2e_steps, "call": "b2ntt_exec_host_batch (pinned host in/out, K steps pipelined over 3 streams)",
                    "blocking_apply_value": e2e_blocking, "blocking_call": "b2ntt_exec_host (AdditiveNTT::apply semantics, one step per call)",
                    "host_path": host_gbs},
            "configs": configs,
            "gpu_launches": Launches.n, "clocks": clocks,
        }
        if world == 1 and not args.no_cpu:
            line["reference_cuda"] = reference_cuda(28 if args.log_n >= 28 else args.log_n)
        print(json.dumps(line), flush=True)
    if world > 1:
        dist.destroy_process_group()


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--log-n", type=int, default=28)
    ap.add_argument("--cpu-log-n", type=int, default=25, help="size of the bounded CPU-baseline sample")
    ap.add_argument("--ref-log-n", type=int, default=24, help="--impl reference: size of the bounded sample each step transforms")
    ap.add_argument("--no-cpu", action="store_true", help="skip the CPU and reference-CUDA baselines")
    ap.add_argument("--no-configs", action="store_true", help="skip the configs block (profiling runs)")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", 0))
    world = int(os.environ.get("WORLD_SIZE", 1))
    local_rank = int(os.environ.get("LOCAL_RANK", 0))
    args.warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    if args.impl == "reference":
        run_reference(args, rank, world)
    else:
        run_ours(args, rank, world, local_rank)


if __name__ == "__main__":
    main()
