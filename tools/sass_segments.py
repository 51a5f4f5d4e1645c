"""Static view of the physics kernel's SASS: splits esb_physics_kernel<8,4> into its functions (at RET) and prints, per
function, the instruction count, the FP64 / MUFU counts and how many MUFU (division / root seeds) sit next to another one
(a pair = two divisions issued side by side, i.e. interleaved dependency chains).
    python tools/sass_segments.py [path/to/libenginesim_b200.so]"""
import os
import re
import subprocess
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    so = sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "engine-sim-cli_b200", "libenginesim_b200.so")
    with tempfile.TemporaryDirectory() as d:
        subprocess.run(["cuobjdump", "-xelf", "physics_kernel", so], cwd=d, check=True, capture_output=True)
        cubin = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
        txt = subprocess.run(["nvdisasm", "-g", os.path.join(d, cubin)], capture_output=True, text=True, check=True).stdout.split("\n")
    start = [i for i, l in enumerate(txt) if l.startswith(".text._ZN4esbk18esb_physics_kernelILi8ELi4E")][0]
    end = [i for i, l in enumerate(txt) if l.startswith(".text.") and i > start][0]
    seg, cur, line = [], [], None
    for l in txt[start:end]:
        m = re.match(r'\s*//## File "([^"]+)", line (\d+)', l)
        if m:
            line = (os.path.basename(m.group(1)), int(m.group(2)))
            continue
        m = re.match(r"\s*/\*([0-9a-f]{4,6})\*/\s+(.*?);", l)
        if m:
            cur.append((m.group(2), line))
            if m.group(2).startswith("RET."):
                seg.append(cur)
                cur = []
    seg.append(cur)
    for k, g in enumerate(seg):
        if len(g) < 100:
            continue
        ops = [(t.split()[1] if t.startswith("@") else t.split()[0]).split(".")[0] for t, _ in g]
        mu = [i for i, o in enumerate(ops) if o == "MUFU"]
        paired = sum(1 for j, i in enumerate(mu) if (j > 0 and i - mu[j - 1] <= 3) or (j + 1 < len(mu) and mu[j + 1] - i <= 3))
        fp64 = sum(1 for o in ops if o in ("DFMA", "DMUL", "DADD", "DSETP"))
        lines = sorted({l[1] for _, l in g if l and l[0] == "physics_core.cuh"})
        print(f"function {k}: {len(g)} instr, FP64 {fp64}, MUFU {len(mu)} ({paired} adjacent to another), LDL {ops.count('LDL')}, STL {ops.count('STL')},"
              f" physics_core.cuh lines {lines[0] if lines else '-'}..{lines[-1] if lines else '-'}")


if __name__ == "__main__":
    main()
