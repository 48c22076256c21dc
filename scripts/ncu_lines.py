"""Per-source-line totals of an .ncu-rep's source page (needs -lineinfo and --import-source on):
executed warp instructions, stall samples and shared-memory wavefronts per file:line, largest first.
usage: python scripts/ncu_lines.py gpurun_out/x.ncu-rep [top_n]"""
import collections
import csv
import io
import subprocess
import sys


def main():
    rep = sys.argv[1]
    top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
    raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass,cuda"],
                         capture_output=True, text=True).stdout
    cur_file, hdr = "", None
    agg = collections.defaultdict(lambda: [0, 0, 0, ""])
    total = [0, 0]
    for row in csv.reader(io.StringIO(raw)):
        if not row:
            continue
        if row[0] == "File Path":
            cur_file = row[1].split("/")[-1]
            continue
        if row[0] == "Line No":
            hdr = row
            continue
        if hdr is None or len(row) < len(hdr) or not row[0].isdigit():
            continue
        d = dict(zip(hdr, row))
        if not d.get("Address"):
            continue  # the CUDA-C view rows; the SASS rows carry the counts and the line they map to
        try:
            ex = int(d.get("Instructions Executed") or 0)
            smp = int(d.get("# Samples") or 0)
            wf = int(d.get("L1 Wavefronts Shared") or 0)
        except ValueError:
            continue
        key = (cur_file, int(row[0]))
        a = agg[key]
        a[0] += ex
        a[1] += smp
        a[2] += wf
        total[0] += ex
        total[1] += smp
    print(f"total executed warp instructions {total[0]}, samples {total[1]}")
    print("file:line | instr | % | samples | % | smem wavefronts")
    for (f, ln), a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
        print(f"{f}:{ln} | {a[0]} | {100 * a[0] / max(total[0], 1):.1f} | {a[1]} | {100 * a[1] / max(total[1], 1):.1f} | {a[2]}")


if __name__ == "__main__":
    main()
