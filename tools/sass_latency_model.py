"""In-order issue model of one warp running alone over the physics kernel's SASS (no branches taken into account: every
function is treated as straight-line code, which the gas exchange functions are).  Latencies are the measured ones
(tools/microbench/fp64_latency.cu on B200: DFMA/DADD/DMUL 8 cycles dependent, 2 cycles issue; MUFU.RCP64H 17; SHFL 12)
plus 4 cycles for the integer / select pipe and 30 for local memory.  It answers one question cheaply, on the CPU:
did a source change shorten the dependency-bound schedule of gasFlow / cellRelax?
    python tools/sass_latency_model.py [disassembly.sass | libenginesim_b200.so]"""
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
FP64 = ("DFMA", "DMUL", "DADD", "DSETP")


def disassemble(so):
    so = os.path.abspath(so)
    with tempfile.TemporaryDirectory() as d:
        subprocess.run(["cuobjdump", "-xelf", "physics_kernel", so], cwd=d, check=True, capture_output=True)
        cubin = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
        return subprocess.run(["nvdisasm", os.path.join(d, cubin)], capture_output=True, text=True, check=True).stdout


def functions(txt):
    """[(name, [instructions])] of the 8-lane kernel: its body, then every out-of-line device function (split at their labels)."""
    txt = txt.split("\n")
    start = [i for i, l in enumerate(txt) if l.startswith(".text._ZN4esbk18esb_physics_kernelILi8ELb0E")][0]
    end = [i for i, l in enumerate(txt) if l.startswith(".text.") and i > start][0]
    seg, cur, name = [], [], "kernel body"
    for l in txt[start:end]:
        m = re.match(r"^\$_ZN4esbk18esb_physics_kernel.*?\$_ZN\d+_INTERNAL_\w+?4esbk(\d+)([A-Za-z0-9_]+)", l)
        if m and l.rstrip().endswith(":"):
            seg.append((name, cur))
            cur, name = [], m.group(2)[:int(m.group(1))]
            continue
        m = re.match(r"^\$__internal_\d+_\$(\w+):", l)
        if m:
            seg.append((name, cur))
            cur, name = [], m.group(1)
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,6})\*/\s+(.*?);", l)
        if m:
            cur.append(m.group(2).strip())
    seg.append((name, cur))
    return seg


def regs(tok, wide):
    out = []
    for m in re.finditer(r"\b(U?R)(\d+)\b|\b(U?P)(\d)\b", tok):
        if m.group(1):
            r = int(m.group(2))
            out.append((m.group(1), r))
            if wide:
                out.append((m.group(1), r + 1))
        else:
            out.append((m.group(3), int(m.group(4))))
    return out


def model(fn, dataflow=False):
    """In-order issue time of the function; dataflow=True: the dependency-only bound (every instruction issues as soon as its
    operands are ready: what a perfect schedule of the same instructions could reach)."""
    ready = {}
    t = 0
    longest = 0
    for ins in fn:
        parts = ins.split(None, 1)
        pred = []
        if parts[0].startswith("@"):
            pred = regs(parts[0].replace("!", ""), False)
            parts = parts[1].split(None, 1)
        op = parts[0]
        base = op.split(".")[0]
        ops = [o.strip() for o in parts[1].split(",")] if len(parts) > 1 else []
        wide = base in FP64 or ".64" in op or ".WIDE" in op
        if base in ("DSETP", "ISETP", "FSETP", "PLOP3"):
            ndst = 2
        elif base in ("STL", "STG", "STS", "BRA", "BSSY", "BSYNC", "RET", "CALL", "BAR", "NOP", "EXIT", "WARPSYNC"):
            ndst = 0
        else:
            ndst = 1
        dst = [r for o in ops[:ndst] for r in regs(o, wide and base != "DSETP")]
        src = [r for o in ops[ndst:] for r in regs(o, wide and base != "MUFU")] + pred
        lat = 8 if base in FP64 else 17 if base == "MUFU" else 12 if base == "SHFL" else 30 if base in ("LDL", "LDG", "LDS", "LDC") else 4
        if base == "DSETP":
            lat = 10
        issue = 2 if base in FP64 else 1
        start = max(([0] if dataflow else [t]) + [ready.get(r, 0) for r in src])
        for r in dst:
            ready[r] = start + lat
        longest = max(longest, start + lat)
        t = start + issue
    return longest if dataflow else t


def main():
    arg = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "engine-sim-cli_b200", "libenginesim_b200.so")
    txt = open(arg).read() if arg.endswith(".sass") else disassemble(arg)
    for name, fn in functions(txt):
        # the fast path of an out-of-line function: up to its first return (slow-path tails and the divergent forms of the shuffles follow it)
        ret = [i for i, ins in enumerate(fn) if ins.startswith("RET")]
        if ret and name != "kernel body":
            fn = fn[:ret[0] + 1]
        if len(fn) >= 300 and name != "kernel body":
            cyc = model(fn)
            nfp = sum(1 for i in fn if (i.split()[1] if i.startswith("@") else i.split()[0]).split(".")[0] in FP64)
            print(f"{name}: {len(fn)} instr, {nfp} FP64 -> {cyc} cycles alone ({cyc / len(fn):.2f} per instr; issue-bound floor {len(fn) + nfp}; "
                  f"dependency-only bound {model(fn, True)})")


if __name__ == "__main__":
    main()
