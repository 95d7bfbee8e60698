#!/usr/bin/env python
"""Which cache-policy instructions survived compilation for sm_100a, per kernel of an object file:
   python profiles/sass_policies.py build/obj/tcif_sweeps_k3.o [regex]
CCTL.E.PF2 = prefetch.global.L2 (pf_l2), LDG.E.EF / STG.E.EF = ld/st.global.cs (__ldcs/__stcs, evict-first),
LDG.E.*.CONSTANT = ld.global.nc (__ldg), *.STRONG.SYS = ld.acquire.sys / st.release.sys (peer-memory halo flags)."""
import re, subprocess, sys
sass = subprocess.run(["cuobjdump", "-sass", sys.argv[1]], capture_output=True, text=True).stdout
pat = re.compile(sys.argv[2]) if len(sys.argv) > 2 else None
name, rows, cnt = None, [], None
keys = ["CCTL.E.PF2", "LDG.E.EF", "STG.E.EF", "CONSTANT", "STRONG.SYS", "LDG", "STG"]
def flush():
    if name and cnt and (not pat or pat.search(name)):
        rows.append((name, dict(cnt)))
for ln in sass.splitlines():
    m = re.search(r"Function : (\S+)", ln)
    if m:
        flush()
        name = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip().split("(")[0].replace("void tcif::", "")
        cnt = {k: 0 for k in keys}
        continue
    if cnt is None:
        continue
    m = re.search(r"\b(CCTL|LDG|STG)[.\w]*", ln)
    if m:
        op = m.group(0)
        for k in keys:
            if k in op:
                cnt[k] += 1
flush()
print(f"{'kernel':60s} " + " ".join(f"{k:>11s}" for k in keys))
for n, c in sorted(rows):
    print(f"{n[:60]:60s} " + " ".join(f"{c[k]:11d}" for k in keys))
