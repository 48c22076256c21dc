"""Summarise an .ncu-rep (read here with `ncu -i`, no GPU needed) into a small markdown table.
usage: python scripts/ncu_summary.py gpurun_out/prof.ncu-rep [title] > profiles/xyz.md"""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("launch__grid_size", "grid"), ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs/thread"),
    ("dram__bytes_read.sum", "dram read"), ("dram__bytes_write.sum", "dram write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram % of peak"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm throughput %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "ALU pipe %"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "FMA pipe %"),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "FP64 pipe %"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "LSU pipe %"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "XU pipe %"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "active threads / instr"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts"),
]


def main():
    rep = sys.argv[1]
    title = sys.argv[2] if len(sys.argv) > 2 else rep
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(raw[raw.index('"ID"'):])))
    hdr, units = rows[0], rows[1]
    print(f"# {title}\n\nsource: `{rep}` (ncu --set full --clock-control none; per-launch, cold-cache, serialised)\n")
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        print(f"## {d['Kernel Name'][:110]}\n\n| metric | value |\n|---|---|")
        for k, label in KEYS:
            if k in d and d[k] not in ("", None):
                print(f"| {label} | {d[k]} {units[hdr.index(k)]} |")
        stalls = []
        for k, v in d.items():
            if "warps_issue_stalled" in k and k.endswith("per_issue_active.ratio") and "not_issued" not in k:
                try:
                    if float(v) > 0.05:
                        stalls.append((float(v), k.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", "")))
                except ValueError:
                    pass
        print("| stall reasons (warps per issue) | " + ", ".join(f"{n} {v:.2f}" for v, n in sorted(stalls, reverse=True)) + " |\n")


if __name__ == "__main__":
    main()
