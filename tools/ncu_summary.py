"""Summarise an .ncu-rep (one kernel launch) into a small JSON for profiles/.  Usage:
python tools/ncu_summary.py gpurun_out/prof.ncu-rep profiles/name.json [steps_in_launch instances]"""
import csv
import json
import subprocess
import sys

KEYS = {
    "gpu__time_duration.sum": "duration",
    "launch__registers_per_thread": "registers_per_thread",
    "launch__grid_size": "grid", "launch__block_size": "block",
    "dram__bytes_read.sum": "dram_read", "dram__bytes_write.sum": "dram_write",
    "smsp__inst_executed.sum": "warp_instructions",
    "smsp__issue_active.avg.pct_of_peak_sustained_active": "issue_active_pct",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "warps_active_pct",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active": "fp64_pipe_active_pct",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active": "fma_pipe_pct",
    "sm__pipe_fmaheavy_cycles_active.avg.pct_of_peak_sustained_active": "fmaheavy_pipe_active_pct",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "sm_throughput_pct",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct",
    "l1tex__t_sector_hit_rate.pct": "l1_hit_pct", "lts__t_sector_hit_rate.pct": "l2_hit_pct",
    "smsp__thread_inst_executed_per_inst_executed.ratio": "threads_per_instruction",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum": "smem_bank_conflicts",
    "smsp__sass_thread_inst_executed_op_ffma_pred_on.sum": "ffma_thread_instr",
    "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum": "dfma_thread_instr",
    "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum": "dmul_thread_instr",
    "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum": "dadd_thread_instr",
}


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, vals = rows[0], rows[1], rows[2]
    d = {"report": rep, "kernel": vals[hdr.index("Kernel Name")] if "Kernel Name" in hdr else None}
    stalls = {}
    for h, u, v in zip(hdr, units, vals):
        if h in KEYS:
            d[KEYS[h]] = {"value": float(v) if v not in ("", "n/a") else None, "unit": u}
        if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
            try:
                stalls[h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]] = round(float(v), 3)
            except ValueError:
                pass
    d["stall_cycles_per_issued_instruction"] = dict(sorted(stalls.items(), key=lambda kv: -kv[1]))
    def to_bytes(x):
        if not x or x["value"] is None:
            return None
        mult = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(x["unit"], 1)
        return x["value"] * mult
    r, w = to_bytes(d.get("dram_read")), to_bytes(d.get("dram_write"))
    if r is not None and w is not None:
        d["dram_bytes_per_launch"] = r + w
    if len(sys.argv) > 4:
        d["steps_in_launch"], d["instances"] = int(sys.argv[3]), int(sys.argv[4])
    json.dump(d, open(out, "w"), indent=1)
    print(json.dumps(d, indent=1)[:1500])


if __name__ == "__main__":
    main()
