"""Warp-stall samples of one kernel in an ncu report, folded onto CUDA source lines.

    python tools/ncu_lines.py report.ncu-rep KERNEL_ID cubin_file function_substring [top]

The report's source page lists SASS addresses; nvdisasm --print-line-info of the matching cubin (cuobjdump -xelf all
libpenroz_b200.so) maps instruction offsets to file:line. Needs a report captured from the same build.
"""
import collections
import csv
import io
import re
import subprocess
import sys

rep, kid, cubin, fn = sys.argv[1:5]
top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--kernel-id", f":::{kid}"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
hi = [i for i, r in enumerate(rows) if "Address" in r and "# Samples" in r][0]
h = rows[hi]
ai, ni, ii = h.index("Address"), h.index("# Samples"), h.index("Instructions Executed")
samples = []
for r in rows[hi + 1:]:
    try:
        samples.append((int(r[ai], 16), int(r[ni]), int(r[ii] or 0)))
    except (ValueError, IndexError):
        pass
base = min(a for a, _, _ in samples)
dis = subprocess.run(["nvdisasm", "--print-line-info", cubin], capture_output=True, text=True).stdout
line_of, cur, infn = {}, None, False
for ln in dis.splitlines():
    m = re.match(r"\s*\.text\.(\S+):", ln)
    if m:
        infn = fn in m.group(1)
        continue
    if not infn:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/", ln)
    if m and cur:
        line_of[int(m.group(1), 16)] = cur
agg = collections.Counter()
inst = collections.Counter()
for a, n, i in samples:
    k = line_of.get(a - base, ("?", 0))
    agg[k] += n
    inst[k] += i
tot = sum(agg.values())
print("samples", tot, "instructions", sum(inst.values()))
for k, n in agg.most_common(top):
    print(f"{n:6d} {100.0 * n / tot:5.1f}%  inst {inst[k]:9d}  {k[0]}:{k[1]}")
