"""profiles/traffic.json from an ncu summary (tools/ncu_summary.py csv of the headline command): per kernel the DRAM bytes and
the executed warp instructions (all pipes / ALU pipe / FMA pipe) per launch, averaged over the captured launches.
usage: python tools/make_traffic.py profiles/<summary>.csv <log_n> '<how it was captured>'"""
import csv, json, sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, body = rows[0], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
acc = {}
for r in body:
    name = r[ix["Kernel Name"]]
    key = "b2::" + name.split("void ")[-1].split("<")[0].split("(")[0].replace("b2::", "")
    a = acc.setdefault(key, {"n": 0, "dram": 0.0, "inst": 0.0, "alu": 0.0, "fma": 0.0})
    a["n"] += 1
    a["dram"] += (float(r[ix["dram__bytes_read.sum"]]) + float(r[ix["dram__bytes_write.sum"]])) * 1e9
    a["inst"] += float(r[ix["smsp__inst_executed.sum"]])
    a["alu"] += float(r[ix["sm__inst_executed_pipe_alu.sum"]])
    a["fma"] += float(r[ix["sm__inst_executed_pipe_fma.sum"]])
out = {"log_n": int(sys.argv[2]), "source": sys.argv[3], "kernels": {}, "algorithmic_bytes_per_launch": 32 << int(sys.argv[2])}
for k, a in acc.items():
    out["kernels"][k] = {"dram_bytes_per_launch": round(a["dram"] / a["n"]), "warp_inst_per_launch": round(a["inst"] / a["n"]),
                         "alu_warp_inst_per_launch": round(a["alu"] / a["n"]), "fma_warp_inst_per_launch": round(a["fma"] / a["n"]),
                         "launches_captured": a["n"]}
json.dump(out, open("profiles/traffic.json", "w"), indent=1)
print(json.dumps(out, indent=1))
