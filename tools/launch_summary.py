"""Summarise an `ncu --metrics gpu__time_duration.sum --csv` launch list into per-kernel totals and shares.
Usage: python tools/launch_summary.py gpurun_out/launches.csv profiles/launches_summary.json "<command that was profiled>" """
import collections
import csv
import json
import sys


def main(src, dst, command):
    rows = [r for r in csv.reader(l for l in open(src) if not l.startswith("=="))]
    hdr, rows = rows[0], rows[1:]
    k, v = hdr.index("Kernel Name"), hdr.index("Metric Value")
    tot = collections.OrderedDict()
    for r in rows:
        name = r[k].split("(")[0]
        t = tot.setdefault(name, [0, 0.0])
        t[0] += 1
        t[1] += float(r[v].replace(",", "")) * 1e-6
    total = sum(t[1] for t in tot.values())
    out = {"command": command,
           "note": "per-launch times under ncu are cold-cache and serialised: compare SHARES with bench.py's kernel_ms, not absolutes",
           "kernels": {n: {"launches": t[0], "total_ms": round(t[1], 3), "share": round(t[1] / total, 4)}
                       for n, t in sorted(tot.items(), key=lambda kv: -kv[1][1])}}
    json.dump(out, open(dst, "w"), indent=1)
    print(json.dumps(out["kernels"], indent=1))


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2], sys.argv[3] if len(sys.argv) > 3 else "")
