"""Static SASS instruction count per source line of one kernel (needs -lineinfo).
usage: python scripts/sass_lines.py file.cubin kernel_substring [top_n]
(cubin: cuobjdump -xelf all foo.o, or the .so)"""
import collections
import re
import subprocess
import sys

cubin, pat = sys.argv[1], sys.argv[2]
top = int(sys.argv[3]) if len(sys.argv) > 3 else 40
dis = subprocess.run(["nvdisasm", "-g", cubin], capture_output=True, text=True).stdout.split("\n")
inside, key, cnt, total = False, "?", collections.Counter(), 0
for line in dis:
    if line.startswith(".text.") or line.lstrip().startswith(".text."):
        inside = pat in line
        continue
    if not inside:
        continue
    m = re.search(r'//## File ".*/([^/"]+)", line (\d+)', line)
    if m:
        key = f"{m.group(1)}:{m.group(2)}"
        continue
    if re.match(r"\s+/\*[0-9a-f]+\*/", line):
        cnt[key] += 1
        total += 1
print("total", total)
for k, v in cnt.most_common(top):
    print(v, k)
