#!/usr/bin/env python
"""Summarise an ncu report by source line: python tools/ncu_lines.py report.ncu-rep [kernel-substring] [top]"""
import collections, csv, subprocess, sys
rep = sys.argv[1]; want = sys.argv[2] if len(sys.argv) > 2 else ""; top = int(sys.argv[3]) if len(sys.argv) > 3 else 30
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "cuda,sass"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = None; cur_file = None; fn = None; seen = set()
agg = collections.defaultdict(lambda: [0, 0]); ops = collections.Counter()
for r in rows:
    if len(r) == 2 and r[0] == "File Path": cur_file = r[1].split('/')[-1]; continue
    if len(r) == 2 and r[0] == "Function Name":
        fn = r[1]; continue
    if r and r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < 10 or want not in (fn or ""): continue
    try:
        ins = int(r[hdr.index("Instructions Executed")] or 0); samp = int(r[hdr.index("# Samples")] or 0)
    except Exception: continue
    key = (cur_file, r[0], r[1].strip()[:100])
    agg[key][0] += ins; agg[key][1] += samp
tot = sum(v[0] for v in agg.values()) or 1; ts = sum(v[1] for v in agg.values()) or 1
print("kernel filter:", want, "| instr", tot, "samples", ts)
for k, v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{100*v[0]/tot:5.1f}% inst {100*v[1]/ts:5.1f}% samp  {k[0]}:{k[1]}  {k[2]}")
