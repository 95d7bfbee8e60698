#!/usr/bin/env python
"""Registers / stack (spill) bytes per kernel of a built object or library:
   python profiles/resource_usage.py build/obj/tcif_sweeps_k3.o [regex]"""
import re, subprocess, sys
out = subprocess.run(["cuobjdump", "--dump-resource-usage", sys.argv[1]], capture_output=True, text=True).stdout
pat = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None
name = None
rows = []
for ln in out.splitlines():
    ln = ln.strip()
    m = re.match(r"Function (\S+):", ln)
    if m:
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip()
        name = name.split("(")[0].replace("void tcif::", "")
        continue
    m = re.match(r"REG:(\d+) STACK:(\d+) SHARED:(\d+)", ln)
    if m and name:
        if not pat or pat.search(name):
            rows.append((name, int(m.group(1)), int(m.group(2)), int(m.group(3))))
        name = None
for r in sorted(rows):
    print(f"{r[1]:4d} regs {r[2]:4d} B stack {r[3]:5d} B smem  {r[0]}")
