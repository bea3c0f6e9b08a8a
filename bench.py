#!/usr/bin/env python3
"""bench.py -- headline benchmark: GF(2^128) additive-NTT elements/s at log_n = 28 on B200.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--log-n L]

A step = one forward additive NTT (b2ntt_exec) of 2^log_n GF(2^128) elements, compact layout,
in place, input resident in HBM.  The data set (4 GiB at log_n = 28) is far larger than the 126 MB
L2, so no flush is needed between steps.  At N > 1 (torchrun, one rank per GPU) every rank runs an
independent replica of the same transform ("replicas only", DESIGN.md section e; weak scaling);
value = N * 2^log_n * K / max-over-ranks device time.

`--impl reference` times the CPU path instead: the reference has no CPU additive NTT, so this is the
oracle port (oracle/antt.c, all host threads) on a bounded sample of the same workload.
"""
import argparse
import json
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


def measured_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "MEASURED_PEAKS.json hbm_gbs"
    except Exception:
        return {"hbm_gbs": 6550.0, "sm_max_mhz": 1965.0}, "fallback of B200_PROFILING.md (6550 GB/s)"


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


def config_dict(args):
    return {"workload": f"additive NTT log_n={args.log_n} GF(2^128) compact layout, log_rate=0, one column per GPU, in place",
            "log_n": args.log_n, "field": "GF(2^128) tower, 4 x u32 limbs", "l2_policy": "data set (4 GiB at log_n=28) >> 126 MB L2; no flush",
            "parallelism": "replicas only (one independent transform per GPU)"}


def cpu_sample(log_h, repeats=1):
    """Time the CPU oracle port (the checker, used here only as the timed CPU baseline) on a bounded
    sample: a GF(2^128) forward transform of 2^log_h elements on all host threads."""
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import oracle as O

    x = np.random.default_rng(1).integers(0, 2**32, size=(1 << log_h) * 4, dtype=np.uint32)
    best = None
    for _ in range(repeats):
        t0 = time.perf_counter()
        O.antt_forward(x, log_h, 0, 4, 0)
        dt = time.perf_counter() - t0
        best = dt if best is None else min(best, dt)
    return {"value": (1 << log_h) / best, "unit": UNIT, "cores": O.lib.orc_num_threads(), "kind": "port",
            "sample": f"one GF(2^128) additive NTT, log_n={log_h}, oracle/antt.c with OpenMP on all host threads "
                      f"({best:.2f} s); the reference has no CPU NTT (SURVEY F5)"}


def run_reference(args, rank, world):
    """--impl reference: the CPU path.  Rank 0 alone works; a step = one log_n=24 sample transform."""
    if rank != 0:
        return
    log_s = min(args.log_n, 24)
    sys.path.insert(0, os.path.join(ROOT, "oracle"))
    import numpy as np
    import oracle as O

    x = np.random.default_rng(1).integers(0, 2**32, size=(1 << log_s) * 4, dtype=np.uint32)
    for _ in range(args.warmup):
        O.antt_forward(x, log_s, 0, 4, 0)
    t0 = time.perf_counter()
    for _ in range(args.steps):
        O.antt_forward(x, log_s, 0, 4, 0)
    dt = time.perf_counter() - t0
    val = args.steps * (1 << log_s) / dt
    cores = O.lib.orc_num_threads()
    sample = f"each step = one GF(2^128) additive NTT of 2^{log_s} elements (bounded sample of the log_n={args.log_n} workload), " \
             f"oracle port on {cores} host threads"
    print(json.dumps({
        "impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
        "warmup": args.warmup, "ms_per_step": dt / args.steps * 1e3, "higher_is_better": True, "scaling": "weak",
        "vs_baseline": None, "dtype": "u32 limbs (GF(2^32) tower arithmetic)", "data": "synthetic",
        "config": config_dict(args),
        "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": "port", "sample": sample},
        "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }), flush=True)


def run_ours(args, rank, world, local_rank):
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
    pass_ms = {}  # kernel name -> [ms of every launch inside the timed region] (CUDA events on the launching stream)
    for _ in range(args.steps):
        for i, (name, t_ms) in enumerate(plan.exec_profile(x, x, stream)):
            pass_ms.setdefault(name, []).append(t_ms)
    e1.record(stream)
    barrier()
    ms = e0.elapsed_time(e1)
    clocks = sampler.finish() if sampler else None
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    ms_max = float(t.item())
    gpu_launches = args.steps * launches_per_step

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
    if world > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    e2e_val = world * n * e2e_steps / float(t[0].item())
    e2e_blocking = world * n / float(t[1].item())
    gpu_launches += (min(e2e_steps, 2) + e2e_steps + 2) * launches_per_step

    if rank == 0:
        peaks, peak_src = measured_peaks()
        peak = float(peaks["hbm_gbs"])
        # dominant kernel = the one with the largest share of the step; its launches were timed live above
        dom = max(pass_ms, key=lambda k: sum(pass_ms[k]))
        launch_ms = sum(pass_ms[dom]) / len(pass_ms[dom])
        ach = BYTES_PER_ELEMENT * n / (launch_ms * 1e-3) / 1e9     # GB/s, algorithmic bytes of one launch (whole data set in + out)
        kernels = {k: {"launches_per_transform": len(v) // args.steps, "avg_launch_ms": sum(v) / len(v),
                       "share_of_step": sum(v) / ms} for k, v in pass_ms.items()}
        traffic = None
        try:
            with open(os.path.join(ROOT, "profiles", "traffic.json")) as f:
                tj = json.load(f)
            if tj.get("log_n") == args.log_n:
                traffic = tj["kernels"][dom]["dram_bytes_per_launch"]
        except Exception:
            pass
        whole = BYTES_PER_ELEMENT * n / (ms / args.steps * 1e-3) / 1e9
        cpu = cpu_sample(args.cpu_log_n) if (world == 1 and not args.no_cpu) else None
        line = {
            "metric": METRIC, "value": world * n * args.steps / (ms_max * 1e-3), "unit": UNIT, "n_gpus": world,
            "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_max / args.steps, "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "u32 limbs (GF(2^32) tower arithmetic, bit-sliced in registers)",
            "data": "synthetic", "config": config_dict(args),
            "roofline": {"bound": "hbm", "kernel": dom, "achieved": ach, "peak": peak, "unit": "GB/s",
                         "frac": ach / peak, "traffic": traffic, "peak_source": peak_src, "kernels": kernels,
                         "launches_per_transform": launches_per_step, "avg_launch_ms": launch_ms,
                         "algorithmic_bytes_per_launch": BYTES_PER_ELEMENT * n,
                         "whole_transform_achieved": whole, "whole_transform_frac": whole / peak},
            "cpu_baseline": cpu,
            "e2e": {"value": e2e_val, "unit": UNIT, "h2d_bytes_per_step": n * 16, "d2h_bytes_per_step": n * 16,
                    "steps": e2e_steps, "call": "b2ntt_exec_host_batch (pinned host in/out, K steps pipelined over 3 streams)",
                    "blocking_apply_value": e2e_blocking, "blocking_call": "b2ntt_exec_host (AdditiveNTT::apply semantics, one step per call)"},
            "gpu_launches": gpu_launches, "clocks": clocks,
        }
        print(json.dumps(line), flush=True)
    plan.close()
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
    ap.add_argument("--no-cpu", action="store_true")
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
