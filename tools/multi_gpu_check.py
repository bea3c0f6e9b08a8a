"""Multi-GPU check and timing (run under torchrun, one rank per GPU):
  torchrun --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/multi_gpu_check.py [--log-n 30]
 1. FourStepNTT parity at log_n = 22 GF(2^128): every rank's block against a single-GPU transform of the whole input;
 2. FourStepNTT timing at --log-n (config 4: one GF(2^128) transform over all GPUs, one NCCL all-to-all);
 3. ColumnShardedRS timing (config 3: 256 columns, log_h = 20, log_rate = 2, GF(2^32), no collective).
Device timing with CUDA events, max over ranks.  Prints one JSON line per item on rank 0."""
import argparse, json, os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
import torch.distributed as dist
import binius_ntt_b200 as b2


def max_ms(ms, dev):
    t = torch.tensor([ms], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--log-n", type=int, default=28)
    ap.add_argument("--iters", type=int, default=3)
    ap.add_argument("--parity-only", action="store_true", help="item 1 only (both exchange paths): the leg tests/test_multi_gpu.py runs")
    args = ap.parse_args()
    rank, world, lr = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(lr)
    dev = torch.device("cuda", lr)
    dist.init_process_group("nccl", device_id=dev)
    # ---- 1. parity of the split against one whole transform --------------------------------------------
    log_s = 22
    g = torch.Generator(device=dev).manual_seed(1234)  # same stream on every rank
    x = torch.randint(-2**31, 2**31 - 1, ((1 << log_s) * 4,), dtype=torch.int32, device=dev, generator=g)
    whole = b2.Plan(128, log_s, 0, device=lr)
    ref = whole.exec(x.clone())
    for fused in (True, False):
        fs = b2.FourStepNTT(log_s, 128, fused=fused)
        mine = b2.FourStepNTT.m_shard(x, world, rank, 4).contiguous().clone()
        out = fs.forward(mine)
        n = fs.local_words
        ok = bool(torch.equal(out, ref[rank * n:(rank + 1) * n]))
        if fused:  # a second transform through the same object: the block is reused behind the barriers
            ok = ok and bool(torch.equal(fs.forward(mine), ref[rank * n:(rank + 1) * n]))
        flags = [None] * world
        dist.all_gather_object(flags, ok)
        if rank == 0:
            print(json.dumps({"check": "four_step_parity", "exchange": "fused peer stores" if fused else "nccl all_to_all", "log_n": log_s, "world": world,
                              "all_blocks_equal": all(flags)}), flush=True)
        fs.close()
    whole.close()
    del x, ref, out, mine
    if args.parity_only:
        dist.destroy_process_group()
        return
    # ---- 2. config 4 timing ----------------------------------------------------------------------------
    fs = b2.FourStepNTT(args.log_n, 128, fused=False)
    buf = torch.randint(-2**31, 2**31 - 1, (fs.local_words,), dtype=torch.int32, device=dev)
    scratch = torch.empty_like(buf)
    stream = torch.cuda.current_stream()
    for _ in range(2):
        fs.forward(buf, scratch)
    torch.cuda.synchronize(); dist.barrier()
    e = [torch.cuda.Event(enable_timing=True) for _ in range(4)]
    tot = [0.0, 0.0, 0.0]
    for _ in range(args.iters):
        e[0].record(); fs.top(buf); e[1].record(); fs.exchange(buf, scratch); e[2].record(); fs.local(scratch); e[3].record()
        torch.cuda.synchronize()
        for i in range(3):
            tot[i] += e[i].elapsed_time(e[i + 1])
    ms = [max_ms(t / args.iters, dev) for t in tot]
    if rank == 0:
        total = sum(ms)
        print(json.dumps({"config": "four_step", "log_n": args.log_n, "field": "GF(2^128)", "world": world, "ms_top": ms[0], "ms_all_to_all": ms[1],
                          "ms_local": ms[2], "ms_total": total, "elements_per_s": (1 << args.log_n) / total * 1e3,
                          "exchange_GBps_per_gpu": fs.local_words * 4 * (world - 1) / world / ms[1] / 1e6 if world > 1 else None}), flush=True)
    fs.close()
    del buf, scratch
    # ---- 3. config 3 timing ----------------------------------------------------------------------------
    rs = b2.ColumnShardedRS(256, 20, 2, field_bits=32)
    xin = torch.randint(-2**31, 2**31 - 1, (rs.in_words,), dtype=torch.int32, device=dev)
    xout = torch.empty(rs.out_words, dtype=torch.int32, device=dev)
    for _ in range(2):
        rs.encode(xin, xout)
    torch.cuda.synchronize(); dist.barrier()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(args.iters):
        rs.encode(xin, xout)
    e1.record(); torch.cuda.synchronize()
    ms3 = max_ms(e0.elapsed_time(e1) / args.iters, dev)
    if rank == 0:
        print(json.dumps({"config": "rs_encode_256col_log20_rate4", "world": world, "columns_per_gpu": rs.count, "ms": ms3,
                          "output_symbols_per_s": 256 * (1 << 22) / ms3 * 1e3, "collective": "none"}), flush=True)
    rs.close()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
