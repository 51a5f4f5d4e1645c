"""Joins an ncu SASS sampling page with nvdisasm's line info for the SAME build and prints where the physics kernel's
warp-stall samples fall: per function segment (split at RET) and, for the kernel body, per enclosing source function.
    ncu -i prof.ncu-rep --page source --csv --print-source sass > sass.csv
    cuobjdump -xelf physics_kernel libenginesim_b200.so && nvdisasm -g physics_kernel.sm_100a.cubin > all.sass
    python tools/ncu_source_profile.py sass.csv all.sass engine-sim-cli_b200/csrc/physics_core.cuh"""
import collections
import csv
import re
import sys


def main():
    sass_csv, disasm, core = sys.argv[1:4]
    lines = open(disasm).read().split("\n")
    start = [i for i, l in enumerate(lines) if l.startswith(".text._ZN4esbk18esb_physics_kernelILi8ELi4E")][0]
    end = [i for i, l in enumerate(lines) if l.startswith(".text.") and i > start][0]
    ins, cur = [], None
    for l in lines[start:end]:
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
        if m:
            cur = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,6})\*/\s+(.*?);", l)
        if m:
            ins.append((m.group(2), cur))
    rows = list(csv.reader(open(sass_csv)))
    h = rows[1]
    idx = {k: i for i, k in enumerate(h)}
    data = [r for r in rows[2:] if len(r) >= len(h)]
    assert len(data) == len(ins), (len(data), len(ins))
    src = open(core).read().split("\n")
    fn_at, cur = [None] * (len(src) + 2), None
    for i, l in enumerate(src, 1):
        m = re.match(r"^(?:template <[^>]*>\s*)?ESB_(?:DEV|MEM)(?:_NOINLINE)?\s+[\w:<>\*& ]+?\s+\*?(\w+)\s*\(", l.strip() if l.startswith("    ESB_MEM") else l)
        if m:
            cur = m.group(1)
        fn_at[i] = cur
    S = sum(int(r[idx["# Samples"]]) for r in data)
    seg, segs, by = 0, collections.Counter(), collections.Counter()
    stall = collections.defaultdict(collections.Counter)
    names = [k for k in h if k.startswith("stall_") and "Not Issued" not in k]
    for (txt, li), r in zip(ins, data):
        s = int(r[idx["# Samples"]])
        segs[seg] += s
        if seg == 0:
            key = fn_at[li[1]] if li and li[0] == "physics_core.cuh" else (li[0] if li else None)
            by[key] += s
            for n in names:
                stall[key][n] += int(r[idx[n]])
        if txt.startswith("RET."):
            seg += 1
    print("samples per function segment (0 = kernel body):", {k: round(v * 100 / S, 1) for k, v in segs.items() if v * 100 / S > 0.3})
    for k, v in by.most_common(14):
        top = ", ".join(f"{n[6:]} {c * 100 // max(v, 1)}%" for n, c in stall[k].most_common(4))
        print(f"  {str(k):28s} {v * 100 / S:5.2f}%   {top}")


if __name__ == "__main__":
    main()
