"""Bucket the SASS page of an ncu report (ncu -i rep --page source --csv --print-source sass > sass.csv) into
regions of N instructions: executed warp-instructions, stall samples and the share of a few stall reasons."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
N = int(sys.argv[2]) if len(sys.argv) > 2 else 500
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
body = rows[2:]
tot_i = sum(float(r[ix["Instructions Executed"]] or 0) for r in body)
tot_s = sum(float(r[ix["# Samples"]] or 0) for r in body)
print(f"total warp-instructions {tot_i:.3e}, samples {tot_s:.0f}")
keys = ["stall_no_inst", "stall_barrier", "stall_wait", "stall_short_sb", "stall_long_sb", "stall_math", "stall_mio", "stall_not_selected", "stall_selected"]
for a in range(0, len(body), N):
    blk = body[a:a + N]
    ins = sum(float(r[ix["Instructions Executed"]] or 0) for r in blk)
    smp = sum(float(r[ix["# Samples"]] or 0) for r in blk)
    ops = {}
    for r in blk:
        op = r[ix["Source"]].split()[0 if not r[ix["Source"]].startswith("@") else 1].split(".")[0]
        ops[op] = ops.get(op, 0) + float(r[ix["Instructions Executed"]] or 0)
    top = sorted(ops.items(), key=lambda kv: -kv[1])[:3]
    st = {k: sum(float(r[ix[k]] or 0) for r in blk) for k in keys}
    print(f"{a:6d} inst {ins / tot_i * 100:5.1f}% samp {smp / tot_s * 100:5.1f}% ", " ".join(f"{k}:{v / ins * 100:.0f}%" for k, v in top if ins),
          " | ", " ".join(f"{k[6:]}:{v / smp * 100:.0f}" for k, v in st.items() if smp and v / smp > 0.04))
