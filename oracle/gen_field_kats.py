#!/usr/bin/env python3
"""Extract the reference's tower-field known-answer vectors (data only) into
tests/golden/field_kats.json.  Run where /root/reference exists; the output is committed.

Sources (paths under /root/reference/src/ulvt/finite_fields/tests/):
  test_fanpaartower.cu:9-120    packed 8/16-bit and compact 32-bit mul / square / inverse
  test_fanpaartower.cu:122-274  bit-sliced 32-bit and 128-bit products (compact images of a, b, result)
  tests.cu:68-91, 172-201       packed H=0,2,5 products and the 128-bit string KAT
"""
import json
import os
import re

REF = "/root/reference/src/ulvt/finite_fields/tests"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def scalar_kats(src):
    out = {"packed_mul": [], "mul32": [], "square32": [], "inverse32": []}
    for m in re.finditer(r"REQUIRE\(mul_binary_tower_32b_simd<(\d)>\((0x[0-9a-fA-F]+),\s*(0x[0-9a-fA-F]+)\)\s*==\s*(0x[0-9a-fA-F]+)\)", src):
        out["packed_mul"].append([int(m.group(1)), int(m.group(2), 16), int(m.group(3), 16), int(m.group(4), 16)])
    for m in re.finditer(r"REQUIRE\(FanPaarTowerField<5>::multiply\((0x[0-9a-fA-F]+),\s*(0x[0-9a-fA-F]+)\)\s*==\s*(0x[0-9a-fA-F]+)\)", src):
        out["mul32"].append([int(m.group(i), 16) for i in (1, 2, 3)])
    for name, key in (("square", "square32"), ("inverse", "inverse32")):
        for m in re.finditer(r"REQUIRE\(FanPaarTowerField<5>::%s\((0x[0-9a-fA-F]+)\)\s*==\s*(0x[0-9a-fA-F]+)\)" % name, src):
            out[key].append([int(m.group(1), 16), int(m.group(2), 16)])
    return out


def array_cases(src):
    """TEST_CASEs that initialise a[], b[] with word lists and REQUIRE results[i] == v."""
    cases = []
    for body in re.split(r"TEST_CASE\(", src)[1:]:
        hm = re.search(r"TEST_TOWER_HEIGHT\s*=\s*(\d)", body)
        arrays = re.findall(r"uint32_t\s+([ab])\[WIDTH\]\s*=\s*\{([^}]*)\}", body)
        res = re.findall(r"REQUIRE\(results?\[(\d+)\]\s*==\s*(0x[0-9a-fA-F]+)\)", body)
        if hm and len(arrays) == 2 and res:
            d = {k: [int(w, 16) for w in re.findall(r"0x[0-9a-fA-F]+", v)] for k, v in arrays}
            cases.append({"h": int(hm.group(1)), "a": d["a"], "b": d["b"], "result": {int(i): int(v, 16) for i, v in res}})
    return cases


def string_case(src):
    m = re.search(r'field_elem_a_str\s*=\s*"0x([0-9a-f]{32})";.*?field_elem_b_str\s*=\s*"0x([0-9a-f]{32})";(.*?)\n}', src, re.S)
    res = re.findall(r"REQUIRE\(result\[(\d)\]\s*==\s*(0x[0-9a-fA-F]+)\)", m.group(3))
    le = lambda s: [int(s[24 - 8 * i:32 - 8 * i], 16) for i in range(4)]  # write_string_to_int_arr: LE words
    return {"h": 7, "a": le(m.group(1)), "b": le(m.group(2)), "result": {int(i): int(v, 16) for i, v in res}}


def main():
    t1 = open(f"{REF}/test_fanpaartower.cu").read()
    t2 = open(f"{REF}/tests.cu").read()
    k = scalar_kats(t1)
    k2 = scalar_kats(t2)
    for key in k:
        k[key] += k2[key]
    k["bitsliced"] = array_cases(t1) + array_cases(t2) + [string_case(t2)]
    k["source"] = "finite_fields/tests/test_fanpaartower.cu, tests.cu (reference known-answer vectors, data only)"
    with open(f"{ROOT}/tests/golden/field_kats.json", "w") as f:
        json.dump(k, f, indent=0, sort_keys=True)
    print({key: len(v) for key, v in k.items() if isinstance(v, list)})


if __name__ == "__main__":
    main()
