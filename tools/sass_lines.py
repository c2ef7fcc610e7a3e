"""Instruction count per source line of one kernel in libsl_gpu.so (nvdisasm -g): where the issue slots go.
    python tools/sass_lines.py <mangled-name-substring> [top]"""
import collections
import os
import re
import subprocess
import sys
import tempfile

so = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "serverless_learn_b200", "libsl_gpu.so")
pat = sys.argv[1]
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
with tempfile.TemporaryDirectory() as d:
    subprocess.run(["cuobjdump", "-xelf", "all", so], cwd=d, check=True, capture_output=True)
    cub = [f for f in os.listdir(d) if f.endswith(".cubin")][0]
    txt = subprocess.run(["nvdisasm", "-g", "-c", os.path.join(d, cub)], capture_output=True, text=True).stdout
cur, on, counts, ops = None, False, collections.Counter(), collections.defaultdict(collections.Counter)
total = 0
for line in txt.splitlines():
    if line.startswith("//--------------------- .text."):
        on = pat in line
        continue
    if not on:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        cur = (os.path.basename(m.group(1)), int(m.group(2)))
        continue
    m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(@!?U?P\d+\s+)?([A-Z0-9_.]+)", line)
    if m:
        counts[cur] += 1
        ops[cur][m.group(2).split(".")[0]] += 1
        total += 1
print("total instructions:", total)
for (f, n), c in counts.most_common(top):
    print(f"{c:6d}  {f}:{n}   " + " ".join(f"{k}x{v}" for k, v in ops[(f, n)].most_common(6)))
