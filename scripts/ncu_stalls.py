"""Per-SASS-instruction stall samples of an .ncu-rep's source page: the instructions with the most samples of one
stall reason (default: all reasons), with the source line they map to.
usage: python scripts/ncu_stalls.py x.ncu-rep [reason|all] [top_n]      reason e.g. long_sb, wait, short_sb, barrier"""
import csv
import io
import subprocess
import sys

rep = sys.argv[1]
reason = sys.argv[2] if len(sys.argv) > 2 else "all"
top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
raw = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"],
                     capture_output=True, text=True).stdout
hdr, rows = None, []
for row in csv.reader(io.StringIO(raw)):
    if row and row[0] == "Address":
        hdr = row
        continue
    if hdr and len(row) >= len(hdr) and row[0].startswith("0x"):
        rows.append(dict(zip(hdr, row)))
col = "# Samples" if reason == "all" else "stall_" + reason
tot = sum(int(r[col] or 0) for r in rows)
print("total", col, tot, "of", sum(int(r["# Samples"] or 0) for r in rows), "samples")
idx = sorted(range(len(rows)), key=lambda i: -int(rows[i][col] or 0))[:top]
for i in sorted(idx):
    r = rows[i]
    prev = rows[i - 1]["Source"].strip() if i else ""
    print(f'{int(r[col] or 0):6d}  {i:5d}  {r["Source"].strip():60s}  <- {prev[:50]}')
