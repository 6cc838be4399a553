#!/usr/bin/env python
"""Opcode histogram + hottest SASS of one kernel in an ncu report: python tools/ncu_sass.py rep kernel-substring [top]"""
import collections, csv, subprocess, sys
rep = sys.argv[1]; want = sys.argv[2]; top = int(sys.argv[3]) if len(sys.argv) > 3 else 25
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = None; fn = None; items = []; done = False
for r in rows:
    if len(r) >= 2 and r[0] in ("Function Name", "Kernel Name"):
        if items and want in (fn or ""): break
        fn = r[1]; continue
    if r and r[0] == "Address": hdr = r; continue
    if hdr is None or want not in (fn or "") or len(r) < len(hdr) - 2: continue
    try: items.append((r[1], int(r[hdr.index("Instructions Executed")] or 0), int(r[hdr.index("# Samples")] or 0)))
    except Exception: pass
tot = sum(i[1] for i in items) or 1; ts = sum(i[2] for i in items) or 1
print(len(items), "SASS instr; executed", tot, "samples", ts)
ops = collections.Counter(); smp = collections.Counter()
for s, e, sa in items:
    t = s.split(); op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
    ops[op] += e; smp[op] += sa
for op, c in ops.most_common(top): print(f"{op:10s} {100*c/tot:5.1f}% inst  {100*smp[op]/ts:5.1f}% samples")
print("--- hottest by samples")
for s, e, sa in sorted(items, key=lambda x: -x[2])[:12]: print(f"{100*sa/ts:5.1f}% {e:9d}  {s[:90]}")
