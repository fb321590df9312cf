"""Turns an .ncu-rep (brought back from the GPU box in gpurun_out/) into the small text
summary committed under profiles/.  Usage: python profiles/summarize_ncu.py <rep> <out.txt> [traffic.json]"""
import csv
import json
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "launch__grid_size", "launch__block_size", "launch__registers_per_thread", "launch__occupancy_limit_registers",
    "launch__occupancy_limit_shared_mem", "launch__waves_per_multiprocessor", "sm__warps_active.avg.per_cycle_active",
    "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
    "smsp__issue_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
    "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    name_i = hdr.index("Kernel Name")
    lines = [f"source: {rep}", ""]
    for r in data:
        lines.append(f"kernel: {r[name_i][:100]}")
        for k in KEYS:
            if k in hdr:
                i = hdr.index(k)
                lines.append(f"  {k:78s} {r[i]:>16s} {units[i]}")
        for i, h in enumerate(hdr):
            if "issue_stalled" in h and h.endswith("per_issue_active.ratio") and float(r[i] or 0) >= 0.05:
                lines.append(f"  stall {h.split('issue_stalled_')[1].split('_per_issue')[0]:72s} {r[i]:>16s}")
        lines.append("")
    open(out, "w").write("\n".join(lines))
    if len(sys.argv) > 3:
        r = data[-1]
        def val(k):
            i = hdr.index(k)
            v = float(r[i].replace(",", ""))
            u = units[i].lower()
            return v * {"byte": 1, "kbyte": 1e3, "mbyte": 1e6, "gbyte": 1e9}[u]
        t = val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
        def num(k):
            return float(r[hdr.index(k)].replace(",", "")) if k in hdr else None
        json.dump({"kernel": r[name_i].split("(")[0], "dram_bytes_per_launch": t, "source": rep,
                   "alu_pipe_busy_pct": num("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
                   "fma_pipe_busy_pct": num("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
                   "issue_slots_busy_pct": num("smsp__issue_active.avg.pct_of_peak_sustained_active"),
                   "warp_instructions": num("smsp__inst_executed.sum"),
                   "ncu_duration_us": num("gpu__time_duration.sum")}, open(sys.argv[3], "w"))
    print("\n".join(lines[:40]))


if __name__ == "__main__":
    main()
