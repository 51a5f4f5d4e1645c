"""Algorithmic FP64 operation count of one physics step (the roofline numerator of bench.py).

Builds the oracle's physics with -fsanitize-coverage=trace-pc, runs a few instances of the bench
grid, and multiplies every basic block's execution count by the FP64 arithmetic instructions the
disassembly shows in that block: add/sub/mul/div/sqrt = 1, fma = 2 (min/max/compare/convert/moves
are not counted).  Output: profiles/flop_counts.json.   Usage: python oracle/flopcount/count_flops.py
"""
import json
import os
import re
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
BUILD = os.path.join(ROOT, "build", "flopcount")

ONE = re.compile(r"^v?(add|sub|mul|div|sqrt)sd$")
FMA = re.compile(r"^vf(n)?m(add|sub)\d+sd$")
PACKED = re.compile(r"^v?(add|sub|mul|div|sqrt)pd$|^vf(n)?m(add|sub)\d+pd$")


def main():
    os.makedirs(BUILD, exist_ok=True)
    exe = os.path.join(BUILD, "flopcount")
    srcs = [os.path.join(ROOT, "oracle", f) for f in ("gas.cpp", "physics.cpp", "solver.cpp", "engine_step.cpp")]
    inc = ["-I" + os.path.join(ROOT, "include"), "-I" + os.path.join(ROOT, "oracle")]
    subprocess.run(["gcc", "-O2", "-c", os.path.join(HERE, "trace.c"), "-o", os.path.join(BUILD, "trace.o")], check=True)
    subprocess.run(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-mfma", "-fno-tree-vectorize", "-fno-tree-slp-vectorize", "-no-pie",
                    "-fno-pie", "-fsanitize-coverage=trace-pc"] + inc + srcs + [os.path.join(HERE, "main.cpp"), os.path.join(BUILD, "trace.o"),
                    "-o", exe], check=True)
    blob = os.path.join(ROOT, "tests", "golden", "gm_ls.params.bin")
    dump = os.path.join(BUILD, "blocks.txt")
    warm, measure, n_inst = 6000, 2000, 16
    steps = int(subprocess.run([exe, blob, str(warm), str(measure), str(n_inst), dump], check=True, capture_output=True, text=True).stdout.strip())
    hits = {}
    for line in open(dump):
        pc, c = line.split()
        hits[int(pc, 16)] = int(c)
    dis = subprocess.run(["objdump", "-d", "--no-show-raw-insn", exe], check=True, capture_output=True, text=True).stdout
    insns = []  # (addr, mnemonic, is_trace_call, starts_function)
    new_fn = False
    for line in dis.splitlines():
        if re.match(r"^[0-9a-f]+ <.*>:$", line):
            new_fn = True
            continue
        m = re.match(r"^\s*([0-9a-f]+):\s+(\S+)(.*)$", line)
        if not m:
            continue
        addr, mn, rest = int(m.group(1), 16), m.group(2), m.group(3)
        insns.append((addr, mn, mn.startswith("call") and "__sanitizer_cov_trace_pc" in rest, new_fn))
        new_fn = False
    totals = {"one": 0, "fma": 0, "packed": 0, "div": 0, "sqrt": 0}
    block_pc, unattributed = None, 0
    for k, (addr, mn, is_call, starts_fn) in enumerate(insns):
        if starts_fn:
            block_pc = None
        if block_pc is not None:
            h = hits.get(block_pc, 0)
            if ONE.match(mn):
                totals["one"] += h
                if "div" in mn:
                    totals["div"] += h
                if "sqrt" in mn:
                    totals["sqrt"] += h
            elif FMA.match(mn):
                totals["fma"] += h
            elif PACKED.match(mn):
                totals["packed"] += h
        elif ONE.match(mn) or FMA.match(mn):
            unattributed += 1
        if is_call and k + 1 < len(insns):
            block_pc = insns[k + 1][0]
    ops = (totals["one"] + 2 * totals["fma"] + 2 * totals["packed"]) / steps
    out = {
        "fp64_ops_per_instance_step_v8": ops,
        "breakdown_per_instance_step": {k: v / steps for k, v in totals.items()},
        "fp64_ops_per_engine_second_v8": ops * 10000,
        "method": "basic-block tracing (-fsanitize-coverage=trace-pc) of oracle/ physics x FP64 arithmetic instructions per block from objdump; "
                  "add/sub/mul/div/sqrt = 1, fma = 2; min/max/compare/convert not counted",
        "sample": f"{n_inst} instances spread over the bench's 64x64 RPM/throttle grid, {measure} steps each after {warm} warm-up steps",
        "static_fp_instructions_before_first_block_of_a_function": unattributed,
    }
    os.makedirs(os.path.join(ROOT, "profiles"), exist_ok=True)
    json.dump(out, open(os.path.join(ROOT, "profiles", "flop_counts.json"), "w"), indent=1)
    print(json.dumps(out, indent=1))


if __name__ == "__main__":
    main()
