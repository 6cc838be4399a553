#!/usr/bin/env python
"""Stall reasons of the hottest SASS region of one kernel launch in an ncu report.

    python tools/ncu_hot.py report.ncu-rep kernel-regex [launch-skip] [min-exec-fraction]

Groups instructions by execution count (a loop body shares one count), prints per group the instruction count,
samples and the stall-reason mix, then the 25 most-sampled instructions with their dominant reason."""
import collections, csv, subprocess, sys
rep, kern = sys.argv[1], sys.argv[2]
skip = sys.argv[3] if len(sys.argv) > 3 else "0"
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass", "--kernel-name",
                      f"regex:{kern}", "--launch-skip", skip, "--launch-count", "1"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr = next(r for r in rows if r and r[0] == "Address")
start = rows.index(hdr)
si, ei = hdr.index("# Samples"), hdr.index("Instructions Executed")
stall = [(i, h) for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
items = []
for r in rows[start + 1:]:
    if len(r) < len(hdr) - 2:
        continue
    try:
        items.append((r[1], int(r[ei] or 0), int(r[si] or 0), {h: int(r[i] or 0) for i, h in stall}))
    except ValueError:
        pass
tot_s = sum(i[2] for i in items) or 1
tot_e = sum(i[1] for i in items) or 1
print(f"{len(items)} SASS instructions, {tot_e} warp-instructions executed, {tot_s} samples")
groups = collections.defaultdict(lambda: [0, 0, collections.Counter()])
for src, e, s, st in items:
    g = groups[e]
    g[0] += 1; g[1] += s; g[2].update(st)
for e, (n, s, st) in sorted(groups.items(), key=lambda kv: -kv[1][1])[:8]:
    mix = ", ".join(f"{k[6:]} {100 * v / max(1, sum(st.values())):.0f}%" for k, v in st.most_common(5))
    print(f"exec {e:7d} x {n:5d} instr ({100 * e * n / tot_e:4.1f}% of executed)  samples {100 * s / tot_s:4.1f}%  | {mix}")
print("--- most-sampled instructions")
for src, e, s, st in sorted(items, key=lambda x: -x[2])[:25]:
    top = st and max(st.items(), key=lambda kv: kv[1])
    print(f"{100 * s / tot_s:4.1f}% exec {e:7d} {top[0][6:] if top else '':12s} {src[:90]}")
