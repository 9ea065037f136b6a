#!/usr/bin/env python
"""Joins an ncu SASS source page (ncu -i X.ncu-rep --page source --csv --print-source sass)
with nvdisasm -g line info of the same cubin, and prints instruction / thread-instruction
counts and stall samples per CUDA source line.  usage: sass_by_line.py sass.csv disasm.txt kernel"""
import collections
import csv
import re
import sys

sass_csv, disasm, kernel = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 50
# ---- nvdisasm: offset -> (file, line)
line_of = {}
cur = None
inside = False
for ln in open(disasm):
    if ln.startswith("//---") and ".text." in ln:
        inside = kernel in ln
        continue
    if not inside:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', ln)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*?);", ln)
    if m:
        line_of[int(m.group(1), 16)] = cur
rows = list(csv.reader(open(sass_csv)))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
base = None
agg = collections.defaultdict(lambda: [0, 0, 0])
tot = [0, 0, 0]
for r in rows[2:]:
    if len(r) < 10 or r[0] == "Address":
        continue
    addr = int(r[0], 16)
    if base is None:
        base = addr
    key = line_of.get(addr - base, ("?", 0))
    n, t, s = int(r[ix["Instructions Executed"]]), int(r[ix["Thread Instructions Executed"]]), int(r[ix["# Samples"]])
    a = agg[key]
    a[0] += n; a[1] += t; a[2] += s
    tot[0] += n; tot[1] += t; tot[2] += s
print(f"total warp inst {tot[0]}  avg active threads {tot[1] / max(tot[0], 1):.2f}  samples {tot[2]}")
for key, (n, t, s) in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{100 * n / tot[0]:5.1f}% inst  thr {t / max(n, 1):5.1f}  {100 * s / max(tot[2], 1):5.1f}% samples  {key[0]}:{key[1]}")
