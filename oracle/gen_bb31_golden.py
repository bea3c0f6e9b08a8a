#!/usr/bin/env python3
"""Extract the reference's BabyBear NTT digests (data only) into tests/golden/bb31_ntt_md5.json.
Source: bb31_ntt_hashes[28][16], /root/reference/src/ulvt/ntt/tests/test_ntt.cu:21-50 (MD5 over asUInt32() of every output,
input = BB31(std::mt19937(0xdeadbeef + log_len)()), generator 137 of the 2^27 subgroup).  Run in the build container."""
import json, os, re

REF = "/root/reference/src/ulvt/ntt/tests/test_ntt.cu"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
src = open(REF).read()
a = src.index("bb31_ntt_hashes[28][16]")
b = src.index("};", a)
rows = []
for m in re.finditer(r"\{0\}|\{((?:\s*0x[0-9a-fA-F]{2}\s*,?\s*){16})\}", src[a:b]):
    rows.append(None if m.group(1) is None else "".join(x[2:].lower() for x in re.findall(r"0x[0-9a-fA-F]{2}", m.group(1))))
assert len(rows) == 28, len(rows)
json.dump({"source": "ntt/tests/test_ntt.cu:21-50", "seed": "mt19937(0xdeadbeef+log_len)", "generator": 137, "log_group_order": 27,
           "md5": {str(i): d for i, d in enumerate(rows) if d}}, open(f"{ROOT}/tests/golden/bb31_ntt_md5.json", "w"), indent=1, sort_keys=True)
print("wrote", len([r for r in rows if r]), "digests")
